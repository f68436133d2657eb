"""``External`` -- the adapter MD engines (TorchMD / OpenMM) call, torchmdnet/calculators.py:21-56.

Same constructor and ``calculate(pos, box)`` contract; the model it drives is the fused CUDA pipeline, so
``device`` must be a CUDA device.  As in the reference, ``box`` is accepted and ignored (calculators.py:50-56);
periodic evaluation is selected on the model (``model.representation_model.distance = OptimizedDistance(...,
box=...)``).
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
    def __init__(self, netfile, embeddings, device="cpu", output_transform=None):
        self.model = load_model(netfile, device=device, derivative=True)
        self.device = device
        self.n_atoms = embeddings.size(1)
        self.embeddings = embeddings.reshape(-1).to(device)
        self.batch = torch.arange(embeddings.size(0), device=device).repeat_interleave(embeddings.size(1))
        self.model.eval()
        # the number of replicas is fixed: tell the head so that no step needs a batch.max() sync
        self.model.output_model.dim_size = int(embeddings.size(0))
        self._n_mol = int(embeddings.size(0))
        if not output_transform:
            self.output_transformer = lambda energy, forces: (energy, forces)
        elif output_transform in tranforms:
            self.output_transformer = tranforms[output_transform]
        else:
            self.output_transformer = eval(output_transform)

    def calculate(self, pos, box):
        pos = pos.to(self.device).type(torch.float32).reshape(-1, 3)
        m = self.model
        mean, std = m._mean_std()
        energy, forces, _ = m.representation_model.energy_forces(self.embeddings, pos, self.batch, self._n_mol,
                                                                 m.output_model, mean, std, compute_forces=True)
        return self.output_transformer(energy.detach(), forces.reshape(-1, self.n_atoms, 3).detach())
