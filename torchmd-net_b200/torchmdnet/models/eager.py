"""Differentiable evaluation of the models for TRAINING: parameter gradients and the
second-order path of a force-matching loss (``autograd.grad(create_graph=True)``, models/model.py:286-298;
caller torchmdnet/module.py:130-179).

The fused kernels produce energies and forces without an autograd graph -- that is the inference / MD hot path.  When
gradients with respect to the PARAMETERS are wanted (``model.train()`` with grad enabled), ``TorchMD_Net.forward`` routes
here instead: the edge list comes from this package's CUDA neighbour op (``OptimizedDistance``; its autograd node has
first AND second derivatives, csrc/torch_shim.cpp), and the model is composed of ordinary differentiable torch ops on
the GPU, applied through the modules' own ``nn.Linear`` / ``nn.Embedding`` / RBF submodules.  Same arithmetic as the
reference's forward (file:line cited per function); no CPU fallback -- CPU tensors are refused by the neighbour op.
"""
from typing import Optional

import torch
import torch.nn.functional as F
from torch import Tensor

_ACT = {"silu": F.silu, "tanh": torch.tanh, "sigmoid": torch.sigmoid, "ssp": lambda x: F.softplus(x) - 0.6931471805599453}


def _segment_sum(src: Tensor, index: Tensor, n: int) -> Tensor:
    return torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device).index_add(0, index, src)


def _edges(rm, pos: Tensor, batch: Tensor):
    """-> (row0, row1, vec, d, self_mask) of the valid pairs: vec = pos[row0] - pos[row1] (+ box shift), d = |vec| with
    d = 0 and vec = 0 on self pairs (the norm of 0 has NaN gradient, models/utils.py:425-438)."""
    ei, d_op, vec = rm.distance(pos, batch)
    ok = ei[0] >= 0                                     # fixed-capacity lists are padded with -1
    if not bool(ok.all()):
        ei, vec = ei[:, ok], vec[ok]
    self_mask = ei[0] == ei[1]
    safe = torch.where(self_mask.unsqueeze(1), torch.ones_like(vec), vec)
    d = torch.where(self_mask, torch.zeros_like(vec[:, 0]), safe.norm(dim=1))
    return ei[0], ei[1], vec * (~self_mask).unsqueeze(1), d, self_mask


# ------------------------------------------------------------------------------------------------ TensorNet
def _skew(v):
    z = torch.zeros_like(v[:, 0])
    return torch.stack([z, -v[:, 2], v[:, 1], v[:, 2], z, -v[:, 0], -v[:, 1], v[:, 0], z], 1).view(-1, 3, 3)


def _decompose(T):
    """tensornet.py:48-54"""
    eye = torch.eye(3, dtype=T.dtype, device=T.device)
    I = T.diagonal(dim1=-2, dim2=-1).mean(-1)[..., None, None] * eye
    A = 0.5 * (T - T.transpose(-2, -1))
    return I, A, 0.5 * (T + T.transpose(-2, -1)) - I


def _tnorm(T):
    return (T ** 2).sum((-2, -1))


def _mix(lin, T):
    """a bias-free channel mix over the H axis of [N,H,3,3] (tensornet.py:317-319)"""
    return lin(T.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)


def tensornet(rm, z: Tensor, pos: Tensor, batch: Tensor) -> Tensor:
    """TensorNet.forward with the unpadded edge semantics (tensornet.py:200-234, 287-326, 376-410) -> x [N,H]"""
    act = _ACT[rm.activation]
    H, n = rm.hidden_channels, z.shape[0]
    i, j, vec, d, self_mask = _edges(rm, pos, batch)
    rbf = rm.distance_expansion(d)
    u = vec / torch.where(self_mask, torch.ones_like(d), d).unsqueeze(1)
    te = rm.tensor_embedding
    C = te.cutoff(d)
    eye = torch.eye(3, dtype=pos.dtype, device=pos.device)
    uu = u[:, :, None] * u[:, None, :]
    Sij = uu - uu.diagonal(dim1=-2, dim2=-1).mean(-1)[:, None, None] * eye
    Z = te.emb(z)
    Zij = te.emb2(torch.cat([Z[i], Z[j]], 1))
    I = _segment_sum((Zij * te.distance_proj1(rbf) * C[:, None])[:, :, None, None] * eye, i, n)
    A = _segment_sum((Zij * te.distance_proj2(rbf) * C[:, None])[:, :, None, None] * _skew(u)[:, None], i, n)
    S = _segment_sum((Zij * te.distance_proj3(rbf) * C[:, None])[:, :, None, None] * Sij[:, None], i, n)
    norm = te.init_norm(_tnorm(I + A + S))
    I, A, S = _mix(te.linears_tensor[0], I), _mix(te.linears_tensor[1], A), _mix(te.linears_tensor[2], S)
    for lin in te.linears_scalar:
        norm = act(lin(norm))
    norm = norm.reshape(n, H, 3)
    X = norm[..., 0, None, None] * I + norm[..., 1, None, None] * A + norm[..., 2, None, None] * S
    for layer in rm.layers:
        f = rbf
        for lin in layer.linears_scalar:
            f = act(lin(f))
        f = (f * layer.cutoff(d)[:, None]).reshape(-1, H, 3)
        X = X / (_tnorm(X) + 1)[..., None, None]
        I, A, S = _decompose(X)
        lt = layer.linears_tensor
        I, A, S = _mix(lt[0], I), _mix(lt[1], A), _mix(lt[2], S)
        Y = I + A + S
        msg = _segment_sum(f[..., 0, None, None] * I[j] + f[..., 1, None, None] * A[j] + f[..., 2, None, None] * S[j], i, n)
        T = msg @ Y + Y @ msg if rm.equivariance_invariance_group == "O(3)" else 2 * (Y @ msg)
        I, A, S = _decompose(T)
        np1 = (_tnorm(I + A + S) + 1)[..., None, None]
        dX = _mix(lt[3], I / np1) + _mix(lt[4], A / np1) + _mix(lt[5], S / np1)
        X = X + dX + dX @ dX
    I, A, S = _decompose(X)
    x = rm.out_norm(torch.cat([_tnorm(I), _tnorm(A), _tnorm(S)], -1))
    return act(rm.linear(x))


# ------------------------------------------------------------------------------------------------ shared pieces
def _neighbor_embedding(ne, z, x, src, tgt, d, rbf, self_mask, n):
    """NeighborEmbedding.forward / message (models/utils.py:73-108): self pairs are dropped"""
    keep = ~self_mask
    W = ne.distance_proj(rbf[keep]) * ne.cutoff(d[keep])[:, None]
    xn = _segment_sum(ne.embedding(z)[src[keep]] * W, tgt[keep], n)
    return ne.combine(torch.cat([x, xn], 1))


# ------------------------------------------------------------------------------------------------ Equivariant Transformer
def equivariant_transformer(rm, z, pos, batch):
    """TorchMD_ET.forward + EquivariantMultiHeadAttention (torchmd_et.py:154-187, 272-347); PyG flow source_to_target:
    source = row 0 of the pair list, target = row 1 -> (x [N,H], vec [N,3,H])"""
    act, attn_act = _ACT[rm.activation], _ACT[rm.attn_activation]
    H, nh, n = rm.hidden_channels, rm.num_heads, z.shape[0]
    hd = H // nh
    src, tgt, vec_e, d, self_mask = _edges(rm, pos, batch)
    rbf = rm.distance_expansion(d)
    dhat = vec_e / torch.where(self_mask, torch.ones_like(d), d).unsqueeze(1)
    x = rm.embedding(z)
    if rm.neighbor_embedding is not None:
        x = _neighbor_embedding(rm.neighbor_embedding, z, x, src, tgt, d, rbf, self_mask, n)
    vec = torch.zeros(n, 3, H, dtype=pos.dtype, device=pos.device)
    for ly in rm.attention_layers:
        xl = ly.layernorm(x)
        q, k = ly.q_proj(xl).view(n, nh, hd), ly.k_proj(xl).view(n, nh, hd)
        v = ly.v_proj(xl).view(n, nh, 3 * hd)
        vec1, vec2, vec3 = torch.split(ly.vec_proj(vec), H, dim=-1)
        vec_dot = (vec1 * vec2).sum(1)
        attn = q[tgt] * k[src]
        if ly.dk_proj is not None:
            attn = attn * act(ly.dk_proj(rbf)).view(-1, nh, hd)
        attn = attn_act(attn.sum(-1)) * ly.cutoff(d)[:, None]
        vj = v[src]
        if ly.dv_proj is not None:
            vj = vj * act(ly.dv_proj(rbf)).view(-1, nh, 3 * hd)
        xm, v1, v2 = torch.split(vj, hd, dim=2)
        vm = vec[src].view(-1, 3, nh, hd) * v1.unsqueeze(1) + v2.unsqueeze(1) * dhat[:, :, None, None]
        xa = _segment_sum(xm * attn.unsqueeze(2), tgt, n).reshape(n, H)
        va = _segment_sum(vm, tgt, n).reshape(n, 3, H)
        o1, o2, o3 = torch.split(ly.o_proj(xa), H, dim=1)
        x = x + vec_dot * o2 + o3
        vec = vec + vec3 * o1.unsqueeze(1) + va
    return rm.out_norm(x), vec


def _gated_block(blk, x, v, act):
    """GatedEquivariantBlock.forward (models/utils.py:496-522)"""
    vb = blk.vec1_proj(v)
    mask = (vb != 0).reshape(vb.size(0), -1).any(1)
    safe = torch.where(mask[:, None, None], vb, torch.ones_like(vb))
    vec1 = torch.where(mask[:, None], safe.norm(dim=-2), torch.zeros_like(vb[:, 0]))
    vec2 = blk.vec2_proj(v)
    h = blk.update_net(torch.cat([x, vec1], -1))
    xo, g = torch.split(h, blk.out_channels, dim=-1)
    return (act(xo) if blk.scalar_activation else xo), g.unsqueeze(1) * vec2


# ------------------------------------------------------------------------------------------------ Transformer, graph network
def transformer(rm, z, pos, batch):
    """TorchMD_T.forward + MultiHeadAttention (torchmd_t.py:144-171, 260-300); self pairs take part in the attention"""
    act, attn_act = _ACT[rm.activation], _ACT[rm.attn_activation]
    H, nh, n = rm.hidden_channels, rm.num_heads, z.shape[0]
    hd = H // nh
    src, tgt, _, d, self_mask = _edges(rm, pos, batch)
    rbf = rm.distance_expansion(d)
    x = rm.embedding(z)
    if rm.neighbor_embedding is not None:
        x = _neighbor_embedding(rm.neighbor_embedding, z, x, src, tgt, d, rbf, self_mask, n)
    for ly in rm.attention_layers:
        xl = ly.layernorm(x)
        q, k, v = (p(xl).view(n, nh, hd) for p in (ly.q_proj, ly.k_proj, ly.v_proj))
        attn = q[tgt] * k[src]
        if ly.dk_proj is not None:
            attn = attn * act(ly.dk_proj(rbf)).view(-1, nh, hd)
        attn = attn_act(attn.sum(-1)) * ly.cutoff(d)[:, None]
        vj = v[src]
        if ly.dv_proj is not None:
            vj = vj * act(ly.dv_proj(rbf)).view(-1, nh, hd)
        x = x + ly.o_proj(_segment_sum(vj * attn.unsqueeze(2), tgt, n).reshape(n, H))
    return rm.out_norm(x)


def graph_network(rm, z, pos, batch):
    """TorchMD_GN.forward + InteractionBlock + CFConv (torchmd_gn.py:148-167, 227-231, 272-290); no self pairs"""
    n = z.shape[0]
    src, tgt, _, d, self_mask = _edges(rm, pos, batch)
    keep = ~self_mask
    src, tgt, d = src[keep], tgt[keep], d[keep]
    rbf = rm.distance_expansion(d)
    x = rm.embedding(z)
    if rm.neighbor_embedding is not None:
        x = _neighbor_embedding(rm.neighbor_embedding, z, x, src, tgt, d, rbf, torch.zeros_like(keep[keep]), n)
    for blk in rm.interactions:
        W = blk.mlp(rbf) * blk.conv.cutoff(d)[:, None]
        agg = _segment_sum(blk.conv.lin1(x)[src] * W, tgt, n)
        x = x + blk.lin(blk.act(blk.conv.lin2(agg)))
    return x


# ------------------------------------------------------------------------------------------------ entry point
def energy(model, z: Tensor, pos: Tensor, batch: Tensor, n_mol: int) -> Tensor:
    """y [n_mol,1] of ``TorchMD_Net`` (models/model.py:255-279: head pre_reduce, * std, atom-wise priors, sum per molecule,
    + mean), differentiable with respect to ``pos`` and every parameter.  Pairwise priors are added by the caller."""
    rm, head = model.representation_model, model.output_model
    kind = rm._op
    if kind == "tensornet":
        x = head.output_network(tensornet(rm, z, pos, batch))
    elif kind == "et":
        x, v = equivariant_transformer(rm, z, pos, batch)
        act = _ACT[rm.activation]
        for blk in head.output_network:
            x, v = _gated_block(blk, x, v, act)
        x = x + v.sum() * 0
    elif kind == "t":
        x = head.output_network(transformer(rm, z, pos, batch))
    elif kind == "gn":
        x = head.output_network(graph_network(rm, z, pos, batch))
    else:
        raise NotImplementedError(kind)
    x = x * model.std
    if model.prior_model is not None:
        from ..priors import Atomref
        for prior in model.prior_model:
            if isinstance(prior, Atomref):
                x = prior.pre_reduce(x, z, pos, batch, None)
    y = _segment_sum(x, batch, n_mol)
    return y + model.mean
