"""B200-native (sm_100a) implementation of TorchMD-NET's energy+force hot path behind the reference's
public names: ``torchmdnet.models.model.{create_model, load_model, TorchMD_Net}``,
``torchmdnet.models.utils.OptimizedDistance``, ``torchmdnet.calculators.External`` and the torch op
``torchmdnet_neighbors::get_neighbor_pairs``.

All compute lives in ``lib/libtmdn.so`` (CUDA kernels behind the C ABI of ``include/tmdn.h``) and is reached
through the thin torch op shim ``lib/libtmdn_torch.so``.  There is no CPU path: importing works anywhere
(so models can be constructed and checkpoints inspected), running requires a CUDA device and the built
libraries, and fails loudly otherwise.
"""
__version__ = "0.1.0"
