"""TEST INFRASTRUCTURE ONLY -- stage-by-stage CPU restatement of the *product's* TensorNet kernel
pipeline (same data layout, same stage boundaries, same hand-derived adjoints as
torchmd-net_b200/csrc/tensornet.cu), in plain torch.

Purpose: (1) the hand-written backward formulas are checked here against torch.autograd of
oracle/model.py (tests/test_staged.py, CPU, fp64) before they are transcribed to CUDA; (2) on the
GPU the intermediate buffers of the CUDA pipeline are compared with the ones produced here.
Never imported by the product.

Layout: node tensors are [N, 9, H] in irreducible components
    c=0      t            I = t*1
    c=1..3   a            A = skew(a) = [[0,-a3,a2],[a3,0,-a1],[-a2,a1,0]]      (tensornet.py:16-34)
    c=4..8   s1..s5       S = [[s1,s3,s4],[s3,s2,s5],[s4,s5,-s1-s2]]            (symmetric traceless)
Edges are CSR by *target* (TensorNet scatters to edge_index[0], gathers from edge_index[1],
tensornet.py:313-315,330-331): row i lists sources j with u = (pos_i - pos_j)/d, d = |pos_i - pos_j|,
including the self loop (u = 0, d = 0).  The edge set is symmetric, so every reverse-edge quantity
needed by the backward pass is recomputed inside row i (no transposed CSR, no atomics).
"""
import math

import torch
import torch.nn.functional as F

GROUP = [0, 1, 1, 1, 2, 2, 2, 2, 2]          # irreps group of component c


# ------------------------------------------------------------------ 3x3 <-> irreps
def to_mat(x):
    """[...,9] -> [...,3,3]"""
    t, a1, a2, a3, s1, s2, s3, s4, s5 = x.unbind(-1)
    rows = [t + s1, s3 - a3, s4 + a2,
            s3 + a3, t + s2, s5 - a1,
            s4 - a2, s5 + a1, t - s1 - s2]
    return torch.stack(rows, -1).reshape(x.shape[:-1] + (3, 3))


def from_mat(T):
    """[...,3,3] -> [...,9]   (decompose_tensor, tensornet.py:48-54)"""
    t = (T[..., 0, 0] + T[..., 1, 1] + T[..., 2, 2]) / 3
    return torch.stack([t,
                        0.5 * (T[..., 2, 1] - T[..., 1, 2]),
                        0.5 * (T[..., 0, 2] - T[..., 2, 0]),
                        0.5 * (T[..., 1, 0] - T[..., 0, 1]),
                        T[..., 0, 0] - t, T[..., 1, 1] - t,
                        0.5 * (T[..., 0, 1] + T[..., 1, 0]),
                        0.5 * (T[..., 0, 2] + T[..., 2, 0]),
                        0.5 * (T[..., 1, 2] + T[..., 2, 1])], -1)


def to_mat_T(G):
    """adjoint of to_mat: dL/dT [...,3,3] -> dL/dx [...,9]"""
    return torch.stack([G[..., 0, 0] + G[..., 1, 1] + G[..., 2, 2],
                        G[..., 2, 1] - G[..., 1, 2],
                        G[..., 0, 2] - G[..., 2, 0],
                        G[..., 1, 0] - G[..., 0, 1],
                        G[..., 0, 0] - G[..., 2, 2], G[..., 1, 1] - G[..., 2, 2],
                        G[..., 0, 1] + G[..., 1, 0],
                        G[..., 0, 2] + G[..., 2, 0],
                        G[..., 1, 2] + G[..., 2, 1]], -1)


def from_mat_T(dx):
    """adjoint of from_mat: dL/dx [...,9] -> dL/dT [...,3,3]"""
    dt, a1, a2, a3, s1, s2, s3, s4, s5 = dx.unbind(-1)
    k = (dt - s1 - s2) / 3
    rows = [k + s1, 0.5 * (s3 - a3), 0.5 * (s4 + a2),
            0.5 * (s3 + a3), k + s2, 0.5 * (s5 - a1),
            0.5 * (s4 - a2), 0.5 * (s5 + a1), k]
    return torch.stack(rows, -1).reshape(dx.shape[:-1] + (3, 3))


def qnorm(x):
    """squared Frobenius norm of to_mat(x) (tensor_norm, tensornet.py:66-67); x [...,9]"""
    t, a1, a2, a3, s1, s2, s3, s4, s5 = x.unbind(-1)
    return 3 * t * t + 2 * (a1 * a1 + a2 * a2 + a3 * a3) + \
        (s1 * s1 + s2 * s2 + (s1 + s2) * (s1 + s2) + 2 * (s3 * s3 + s4 * s4 + s5 * s5))


def qnorm_parts(x):
    t, a1, a2, a3, s1, s2, s3, s4, s5 = x.unbind(-1)
    return (3 * t * t, 2 * (a1 * a1 + a2 * a2 + a3 * a3),
            s1 * s1 + s2 * s2 + (s1 + s2) * (s1 + s2) + 2 * (s3 * s3 + s4 * s4 + s5 * s5))


def qnorm_grad(x):
    t, a1, a2, a3, s1, s2, s3, s4, s5 = x.unbind(-1)
    return torch.stack([6 * t, 4 * a1, 4 * a2, 4 * a3, 4 * s1 + 2 * s2, 4 * s2 + 2 * s1, 4 * s3, 4 * s4, 4 * s5], -1)


# ------------------------------------------------------------------ radial functions with d/dd
def act_fn(name, x):
    if name == "silu":
        s = torch.sigmoid(x)
        return x * s, s * (1 + x * (1 - s))
    if name == "tanh":
        y = torch.tanh(x)
        return y, 1 - y * y
    if name == "sigmoid":
        s = torch.sigmoid(x)
        return s, s * (1 - s)
    if name == "ssp":
        return F.softplus(x) - math.log(2.0), torch.sigmoid(x)
    raise ValueError(name)


def cutoff_and_grad(d, lo, hi):
    """CosineCutoff (models/utils.py:368-390) and dC/dd."""
    if lo > 0:
        arg = math.pi * (2 * (d - lo) / (hi - lo) + 1.0)
        m = ((d < hi) & (d > lo)).to(d.dtype)
        return 0.5 * (torch.cos(arg) + 1) * m, -0.5 * torch.sin(arg) * (2 * math.pi / (hi - lo)) * m
    m = (d < hi).to(d.dtype)
    return 0.5 * (torch.cos(d * math.pi / hi) + 1) * m, -0.5 * torch.sin(d * math.pi / hi) * (math.pi / hi) * m


def rbf_and_grad(d, hp, p0, p1):
    """[E] -> rbf[E,K], drbf/dd[E,K].  expnorm: p0=means p1=betas (models/utils.py:339-344); gauss: p0=offset p1=coeff."""
    lo, hi = hp["cutoff_lower"], hp["cutoff_upper"]
    dd = d.unsqueeze(-1)
    if hp["rbf_type"] == "expnorm":
        alpha = 5.0 / (hi - lo)
        c0, dc0 = cutoff_and_grad(dd, 0.0, hi)
        u = torch.exp(alpha * (lo - dd))
        g = torch.exp(-p1 * (u - p0) ** 2)
        return c0 * g, dc0 * g + c0 * g * (2 * alpha) * p1 * u * (u - p0)
    g = torch.exp(p1 * (dd - p0) ** 2)
    return g, 2 * p1 * (dd - p0) * g


def filter_mlp(d, hp, rbfp, layers, act, final_act):
    """f(d) = [act∘]Lin_n(...act(Lin_1(rbf(d)))) * C(d)  and df/dd, by forward-mode tangents.
    `layers` = [(W,b), ...] with W[out,in]."""
    lo, hi = hp["cutoff_lower"], hp["cutoff_upper"]
    h, t = rbf_and_grad(d, hp, *rbfp)
    for k, (W, b) in enumerate(layers):
        h = F.linear(h, W, b)
        t = F.linear(t, W)
        if k < len(layers) - 1 or final_act:
            h, g = act_fn(act, h)
            t = t * g
    c, dc = cutoff_and_grad(d, lo, hi)
    return h * c[:, None], t * c[:, None] + h * dc[:, None]


def edge_geom(u):
    """geom_c(u) [E,9] and d geom_c / d u [E,9,3] for the embedding (tensornet.py:38-44, :16-34)."""
    ux, uy, uz = u.unbind(-1)
    n2 = (ux * ux + uy * uy + uz * uz) / 3
    one, zero = torch.ones_like(ux), torch.zeros_like(ux)
    g = torch.stack([one, ux, uy, uz, ux * ux - n2, uy * uy - n2, ux * uy, ux * uz, uy * uz], -1)
    tw = 2.0 / 3.0
    dg = torch.stack([
        torch.stack([zero, zero, zero], -1),
        torch.stack([one, zero, zero], -1), torch.stack([zero, one, zero], -1), torch.stack([zero, zero, one], -1),
        torch.stack([2 * ux - tw * ux, -tw * uy, -tw * uz], -1),
        torch.stack([-tw * ux, 2 * uy - tw * uy, -tw * uz], -1),
        torch.stack([uy, ux, zero], -1), torch.stack([uz, zero, ux], -1), torch.stack([zero, uz, uy], -1)], 1)
    return g, dg


# ------------------------------------------------------------------ packed weights
def pack_tensornet(sd, hp, pre="representation_model."):
    """What the host side precomputes once per model (torchmdnet/models/tensornet.py in the product)."""
    H = hp["embedding_dimension"]
    te = pre + "tensor_embedding."
    W = {}
    emb = sd[te + "emb.weight"]
    w2, b2 = sd[te + "emb2.weight"], sd[te + "emb2.bias"]
    W["P"] = emb @ w2[:, :H].T + b2          # Z_ij = P[z_i] + Q[z_j]     (tensornet.py:310-312)
    W["Q"] = emb @ w2[:, H:].T
    W["emb_filter"] = [(torch.cat([sd[te + "distance_proj%d.weight" % k] for k in (1, 2, 3)]),
                        torch.cat([sd[te + "distance_proj%d.bias" % k] for k in (1, 2, 3)]))]
    perm = torch.arange(3 * H).view(H, 3).t().reshape(-1)   # row g*H+h <- reference row h*3+g
    W["emb_mix"] = [sd[te + "linears_tensor.%d.weight" % k] for k in range(3)]
    W["emb_ln"] = (sd[te + "init_norm.weight"], sd[te + "init_norm.bias"])
    W["emb_mlp"] = [(sd[te + "linears_scalar.0.weight"], sd[te + "linears_scalar.0.bias"]),
                    (sd[te + "linears_scalar.1.weight"][perm], sd[te + "linears_scalar.1.bias"][perm])]
    W["layers"] = []
    for l in range(hp["num_layers"]):
        lp = pre + "layers.%d." % l
        flt = [(sd[lp + "linears_scalar.%d.weight" % k], sd[lp + "linears_scalar.%d.bias" % k]) for k in range(3)]
        flt[2] = (flt[2][0][perm], flt[2][1][perm])
        W["layers"].append(dict(filter=flt, mix=[sd[lp + "linears_tensor.%d.weight" % k] for k in range(6)]))
    W["out_ln"] = (sd[pre + "out_norm.weight"], sd[pre + "out_norm.bias"])
    W["out_lin"] = (sd[pre + "linear.weight"], sd[pre + "linear.bias"])
    W["head"] = [(sd["output_model.output_network.0.weight"], sd["output_model.output_network.0.bias"]),
                 (sd["output_model.output_network.2.weight"], sd["output_model.output_network.2.bias"])]
    if hp["rbf_type"] == "expnorm":
        W["rbf"] = (sd[pre + "distance_expansion.means"], sd[pre + "distance_expansion.betas"])
    else:
        W["rbf"] = (sd[pre + "distance_expansion.offset"], sd[pre + "distance_expansion.coeff"])
    W["std"] = sd.get("std", torch.tensor(1.0))
    W["mean"] = sd.get("mean", torch.tensor(0.0))
    return W


def mix(x, Ws):
    """grouped channel mix: out[n,c,:] = W_{g(c)} x[n,c,:]"""
    return torch.stack([F.linear(x[:, c], Ws[GROUP[c]]) for c in range(9)], 1)


def mix_T(dx, Ws):
    return torch.stack([dx[:, c] @ Ws[GROUP[c]] for c in range(9)], 1)


def layer_norm_fwd(x, g, b, eps=1e-5):
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    rstd = 1 / torch.sqrt(var + eps)
    xh = (x - mu) * rstd
    return xh * g + b, (xh, rstd)


def layer_norm_bwd(dy, g, saved):
    xh, rstd = saved
    dg = dy * g
    return (dg - dg.mean(-1, keepdim=True) - xh * (dg * xh).mean(-1, keepdim=True)) * rstd


# ------------------------------------------------------------------ the staged pipeline
def csr_from_edges(ei, n):
    """ei[0]=target i, ei[1]=source j  ->  row_ptr, col, perm (edges sorted by (i, j))."""
    key = ei[0] * n + ei[1]
    perm = torch.argsort(key, stable=True)
    tgt, col = ei[0][perm], ei[1][perm]
    row_ptr = torch.zeros(n + 1, dtype=torch.long)
    row_ptr[1:] = torch.cumsum(torch.bincount(tgt, minlength=n), 0)
    return row_ptr, tgt, col, perm


def energy_forces_staged(sd, hp, z, pos, batch, ei, shift, n_mol, keep=None):
    """Forward through the product's stages, then the hand-derived backward.  Returns (y[M,1], forces[N,3]).
    `keep` (dict) receives every intermediate buffer under the names the CUDA workspace uses."""
    keep = {} if keep is None else keep
    W = pack_tensornet(sd, hp)
    act = hp["activation"]
    H, n = hp["embedding_dimension"], z.shape[0]
    o3 = hp.get("equivariance_invariance_group", "O(3)") == "O(3)"
    G = torch.tensor(GROUP)
    row_ptr, tgt, col, perm = csr_from_edges(ei, n)
    vec = pos[tgt] - pos[col] + shift[perm]
    selfm = tgt == col
    d = torch.where(selfm, torch.zeros_like(vec[:, 0]), vec.norm(dim=1))
    u = torch.where(selfm[:, None], torch.zeros_like(vec), vec / torch.where(selfm, torch.ones_like(d), d)[:, None])
    keep.update(row_ptr=row_ptr, col=col, u=u, d=d)

    # ---------------- forward
    # E1: embedding edge kernel
    Fe, dFe = filter_mlp(d, hp, W["rbf"], W["emb_filter"], act, final_act=False)       # [E,3H] (g*H+h)
    Fe, dFe = Fe.view(-1, 3, H), dFe.view(-1, 3, H)
    geom, dgeom = edge_geom(u)
    Zij = W["P"][z[tgt]] + W["Q"][z[col]]
    msg = Zij[:, None, :] * Fe[:, G, :] * geom[:, :, None]
    ias = torch.zeros(n, 9, H, dtype=pos.dtype).index_add(0, tgt, msg)
    nrm = qnorm(ias.transpose(1, 2))                                                    # [N,H]
    nh, ln0 = layer_norm_fwd(nrm, *W["emb_ln"])
    h1 = F.linear(nh, *W["emb_mlp"][0]); a1, g1 = act_fn(act, h1)
    h2 = F.linear(a1, *W["emb_mlp"][1]); a2, g2 = act_fn(act, h2)
    fias = a2.view(n, 3, H)
    iasm = mix(ias, W["emb_mix"])
    X = fias[:, G, :] * iasm
    keep.update(ias=ias, nh=nh, fias=fias, X0=X)
    saved = []
    for l, L in enumerate(W["layers"]):
        Fl, dFl = filter_mlp(d, hp, W["rbf"], L["filter"], act, final_act=True)
        Fl, dFl = Fl.view(-1, 3, H), dFl.view(-1, 3, H)
        qx = qnorm(X.transpose(1, 2))                                                   # [N,H]
        Xh = X / (qx + 1)[:, None, :]
        Y = mix(Xh, L["mix"][:3])
        Mm = torch.zeros_like(Y).index_add(0, tgt, Fl[:, G, :] * Y[col])
        Ym, Mmat = to_mat(Y.transpose(1, 2)), to_mat(Mm.transpose(1, 2))                # [N,H,3,3]
        T = Mmat @ Ym + Ym @ Mmat if o3 else 2 * (Ym @ Mmat)
        tau = from_mat(T)                                                               # [N,H,9]
        qt = qnorm(tau)
        tauh = (tau / (qt + 1)[..., None]).transpose(1, 2)                              # [N,9,H]
        dX = mix(tauh, L["mix"][3:])
        Dm = to_mat(dX.transpose(1, 2))
        Xn = Xh + dX + from_mat(Dm @ Dm).transpose(1, 2)
        saved.append(dict(X=X, qx=qx, Xh=Xh, Y=Y, M=Mm, tau=tau, qt=qt, dX=dX, F=Fl, dF=dFl))
        keep.update({"Y%d" % l: Y, "M%d" % l: Mm, "tauh%d" % l: tauh, "X%d" % (l + 1): Xn})
        X = Xn
    pI, pA, pS = qnorm_parts(X.transpose(1, 2))
    xo = torch.cat([pI, pA, pS], -1)                                                    # [N,3H]
    xl, ln1 = layer_norm_fwd(xo, *W["out_ln"])
    h3 = F.linear(xl, *W["out_lin"]); a3, g3 = act_fn(act, h3)
    h4 = F.linear(a3, *W["head"][0]); a4, g4 = act_fn(act, h4)
    e_atom = F.linear(a4, *W["head"][1]) * W["std"]
    y = torch.zeros(n_mol, 1, dtype=pos.dtype).index_add(0, batch, e_atom) + W["mean"]
    keep.update(e_atom=e_atom, xl=xl)

    # ---------------- backward (dL/dy = 1)
    d4 = (torch.ones(n, 1, dtype=pos.dtype) * W["std"]) @ W["head"][1][0] * g4
    d3 = (d4 @ W["head"][0][0]) * g3
    dxl = d3 @ W["out_lin"][0]
    dxo = layer_norm_bwd(dxl, W["out_ln"][0], ln1)                                      # [N,3H]
    # d/dx of the three partial norms: qnorm_grad restricted to each group
    gq = qnorm_grad(X.transpose(1, 2)).transpose(1, 2)                                  # [N,9,H]
    dXc = gq * dxo.view(n, 3, H)[:, G, :]
    gpos = torch.zeros(n, 3, dtype=pos.dtype)
    Gn = torch.zeros(col.shape[0], dtype=pos.dtype)   # per-edge dE/dd_e summed with its reverse edge (debug only)
    for l in reversed(range(len(W["layers"]))):
        L, S = W["layers"][l], saved[l]
        # Xn = Xh + dX + from_mat(Dm@Dm)
        Gs = from_mat_T(dXc.transpose(1, 2))                                            # [N,H,3,3]
        Dm = to_mat(S["dX"].transpose(1, 2))
        ddX = dXc + to_mat_T(Gs @ Dm.transpose(-1, -2) + Dm.transpose(-1, -2) @ Gs).transpose(1, 2)
        dXh = dXc.clone()
        dtauh = mix_T(ddX, L["mix"][3:])                                                # [N,9,H]
        # tauh = tau/(qt+1)
        tau, qt = S["tau"], S["qt"]
        dth = dtauh.transpose(1, 2)                                                     # [N,H,9]
        dtau = dth / (qt + 1)[..., None] - ((dth * tau).sum(-1) / (qt + 1) ** 2)[..., None] * qnorm_grad(tau)
        GT = from_mat_T(dtau)
        Ym, Mmat = to_mat(S["Y"].transpose(1, 2)), to_mat(S["M"].transpose(1, 2))
        if o3:
            dMmat = GT @ Ym.transpose(-1, -2) + Ym.transpose(-1, -2) @ GT
            dYmat = Mmat.transpose(-1, -2) @ GT + GT @ Mmat.transpose(-1, -2)
        else:
            dMmat = 2 * (Ym.transpose(-1, -2) @ GT)
            dYmat = 2 * (GT @ Mmat.transpose(-1, -2))
        dM = to_mat_T(dMmat).transpose(1, 2)                                            # [N,9,H]
        dY = to_mat_T(dYmat).transpose(1, 2)
        # E3: message-passing backward inside row i (symmetric edge set):
        #   dY_i += sum_e F_e * dM_j ;  gpos_i += u_e * sum_h,c dF_e[g(c)] (dM_c[i] Y_c[j] + dM_c[j] Y_c[i])
        Fg, dFg = S["F"][:, G, :], S["dF"][:, G, :]
        dY = dY + torch.zeros_like(dY).index_add(0, tgt, Fg * dM[col])
        ge = (dFg * (dM[tgt] * S["Y"][col] + dM[col] * S["Y"][tgt])).sum((1, 2))
        gpos = gpos.index_add(0, tgt, u * ge[:, None])
        dXh = dXh + mix_T(dY, L["mix"][:3])
        # Xh = X/(qx+1)
        Xl, qx = S["X"], S["qx"]
        dXc = dXh / (qx + 1)[:, None, :] - ((dXh * Xl).sum(1) / (qx + 1) ** 2)[:, None, :] * \
            qnorm_grad(Xl.transpose(1, 2)).transpose(1, 2)
    # embedding: X = fias[g] * iasm
    dfias = torch.stack([(dXc * iasm)[:, [c for c in range(9) if GROUP[c] == g]].sum(1) for g in range(3)], 1)
    diasm = dXc * fias[:, G, :]
    dh2 = dfias.reshape(n, 3 * H) * g2
    dh1 = (dh2 @ W["emb_mlp"][1][0]) * g1
    dnh = dh1 @ W["emb_mlp"][0][0]
    dnrm = layer_norm_bwd(dnh, W["emb_ln"][0], ln0)                                     # [N,H]
    Dias = mix_T(diasm, W["emb_mix"]) + dnrm[:, None, :] * qnorm_grad(ias.transpose(1, 2)).transpose(1, 2)
    keep.update(Dias=Dias)
    # E4: embedding edge backward inside row i: edge e=(i,j) and its reverse e'=(j,i) (u' = -u, same d)
    Zji = W["P"][z[col]] + W["Q"][z[tgt]]
    geom_r, dgeom_r = edge_geom(-u)
    A_e = Dias[tgt] * Zij[:, None, :]                                                   # [E,9,H]
    A_r = Dias[col] * Zji[:, None, :]
    gd = (A_e * dFe[:, G, :] * geom[:, :, None]).sum((1, 2)) + (A_r * dFe[:, G, :] * geom_r[:, :, None]).sum((1, 2))
    wu_e = (A_e * Fe[:, G, :]).sum(2)                                                   # [E,9]
    wu_r = (A_r * Fe[:, G, :]).sum(2)
    gu_e = (wu_e[:, :, None] * dgeom).sum(1)                                            # dL/du_e   [E,3]
    gu_r = (wu_r[:, :, None] * dgeom_r).sum(1)                                          # dL/du_e'
    inv_d = torch.where(selfm, torch.zeros_like(d), 1 / torch.where(selfm, torch.ones_like(d), d))
    # dL/dvec_e = gd_e u + (gu_e - (gu_e.u)u)/d ;  atom i receives  +dL/dvec_e  -dL/dvec_e'
    perp_e = (gu_e - (gu_e * u).sum(1, keepdim=True) * u) * inv_d[:, None]
    perp_r = (gu_r - (gu_r * u).sum(1, keepdim=True) * u) * inv_d[:, None]
    gpos = gpos.index_add(0, tgt, u * gd[:, None] + perp_e - perp_r)
    return y, -gpos
