"""``External`` -- the adapter MD engines (TorchMD / OpenMM) call, torchmdnet/calculators.py:21-56.

Same constructor and ``calculate(pos, box)`` contract; the model it drives is the fused CUDA pipeline, so
``device`` must be a CUDA device.  Because the atom types and the number of replicas are fixed at construction,
the whole step (neighbour list, forward, analytic backward) is captured once into a CUDA graph after two eager
warm-up calls and replayed afterwards (``use_cuda_graph=False`` keeps eager launches).

``box``: the reference accepts and ignores it (calculators.py:50-56) -- its models never pass a box to their distance
module.  Here a box that is given is USED: ``None`` or an all-zero box evaluates open boundaries as
before; a [3,3] (or [R,3,3]: the first replica's box is used for all, or per-axis lengths [3]) box switches the model's neighbour search to periodic boundaries,
with the cell list when the box is rectangular and at least three cutoffs wide, else the brute strategy (triclinic
boxes, small boxes).  The box lives on the host (the op takes it by value); a change of box re-captures the graph.
"""
import torch

from .models.model import load_model

# preset unit transforms (calculators.py:4-18)
tranforms = {
    "eV/A -> kcal/mol/A": lambda energy, forces: (energy * 23.0609, forces * 23.0609),
    "Hartree/Bohr -> kcal/mol/A": lambda energy, forces: (energy * 627.509, forces * 627.509 / 0.529177),
    "Hartree/A -> kcal/mol/A": lambda energy, forces: (energy * 627.509, forces * 627.509),
}


class External:
    def __init__(self, netfile, embeddings, device="cpu", output_transform=None, use_cuda_graph=True):
        self.model = load_model(netfile, device=device, derivative=True)
        self.device = device
        self.n_atoms = embeddings.size(1)
        self.embeddings = embeddings.reshape(-1).to(device)
        self.batch = torch.arange(embeddings.size(0), device=device).repeat_interleave(embeddings.size(1))
        self.model.eval()
        # the number of replicas is fixed: tell the head so that no step needs a batch.max() sync
        self.model.output_model.dim_size = int(embeddings.size(0))
        self._n_mol = int(embeddings.size(0))
        self.use_cuda_graph = use_cuda_graph
        self._calls = 0
        self._graph = None
        self._static_pos = None
        self._static_out = None
        self._box_key = None
        self._box_tag = None
        if not output_transform:
            self.output_transformer = lambda energy, forces: (energy, forces)
        elif output_transform in tranforms:
            self.output_transformer = tranforms[output_transform]
        else:
            self.output_transformer = eval(output_transform)

    def _evaluate(self, pos):
        m = self.model
        mean, std = m._mean_std()
        energy, forces, _ = m.representation_model.energy_forces(self.embeddings, pos, self.batch, self._n_mol,
                                                                 m.output_model, mean, std, compute_forces=True)
        return m.apply_priors(energy, self.embeddings, self.batch, self._n_mol, pos, forces), forces

    def invalidate_graph(self):
        """Call after changing the model (weights, filter mode, distance module) to re-capture."""
        self._graph, self._calls = None, 0

    def _use_box(self, box):
        """Route ``box`` into the model's distance module; -> True if the evaluation setup changed."""
        if box is None:
            key = None
        else:
            if isinstance(box, torch.Tensor):       # same tensor, unchanged since the last call: nothing to do (no sync)
                tag = (id(box), box._version)
                if tag == self._box_tag:
                    return False
                self._box_tag = tag
            b = torch.as_tensor(box).detach().to("cpu", torch.float64)
            if b.numel() == 3:
                b = torch.diag(b.reshape(3))
            b = b.reshape(-1, 3, 3)[0]
            key = None if float(b.abs().max()) == 0.0 else tuple(b.reshape(-1).tolist())
        if key == self._box_key:
            return False
        from .models.utils import OptimizedDistance
        rm = self.model.representation_model
        old = rm.distance
        kw = dict(max_num_pairs=old.max_num_pairs, return_vecs=True, loop=True, check_errors=old.check_errors,
                  resize_to_fit=old.resize_to_fit, long_edge_index=True)
        if key is None:
            rm.distance = OptimizedDistance(old.cutoff_lower, old.cutoff_upper, **kw)
        else:
            b = torch.tensor(key, dtype=torch.float64).reshape(3, 3)
            diag = bool((b - torch.diag(torch.diagonal(b))).abs().max() == 0)
            cell = diag and float(torch.diagonal(b).min()) >= 3.0 * float(old.cutoff_upper)
            rm.distance = OptimizedDistance(old.cutoff_lower, old.cutoff_upper, box=b.to(torch.float32),
                                            strategy="cell" if cell else "brute", **kw)
        self._box_key = key
        return True

    def calculate(self, pos, box):
        if self._use_box(box):
            self.invalidate_graph()
        self._calls += 1
        if not self.use_cuda_graph or self._calls <= 2:
            pos = pos.to(self.device, non_blocking=True).type(torch.float32).reshape(-1, 3)
            energy, forces = self._evaluate(pos)
        else:
            if self._graph is None:
                self._static_pos = torch.empty(self.embeddings.numel(), 3, device=self.device)
                self._static_pos.copy_(pos.reshape(-1, 3))
                torch.cuda.synchronize()
                self._graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._graph):
                    self._static_out = self._evaluate(self._static_pos)
            self._static_pos.copy_(pos.reshape(-1, 3), non_blocking=True)     # H2D (or D2D) + dtype cast
            self._graph.replay()
            energy, forces = (t.clone() for t in self._static_out)
        return self.output_transformer(energy.detach(), forces.reshape(-1, self.n_atoms, 3).detach())
