"""TEST INFRASTRUCTURE ONLY -- stage-by-stage CPU restatement of the *product's* Equivariant Transformer kernel
pipeline (same data layout, stage boundaries and hand-derived adjoints as torchmd-net_b200/csrc/et.cu), in plain
torch.  tests/test_staged.py checks it against torch.autograd of oracle/model.py in fp64 before the formulas are
transcribed to CUDA; the GPU tests compare intermediate buffers with it.  Never imported by the product.

Layout: x [N,H]; vec [N,3,H]; v and dv are *part-major* [.,3,H] (part 0 = scalar message, 1 = vec scale,
2 = direction scale) with channel h = head*hd + c, i.e. the reference's per-head interleave
[x | v1 | v2] (torchmd_et.py:327) is undone once when the weights are packed.
Edges are CSR by *target* i (reference: index = edge_index[1]; sources j = edge_index[0]); the stored unit vector
is u = (pos_i - pos_j)/d, so the reference's d_ij = (pos_j - pos_i)/d = -u.  The edge set is symmetric, so row i
also evaluates the reverse edge (j <- i) of every entry and collects the source-side adjoints (k, v, vec of atom i)
without a transposed graph or atomics.
"""
import torch
import torch.nn.functional as F

from .staged_tensornet import act_fn, csr_from_edges, cutoff_and_grad, layer_norm_bwd, layer_norm_fwd, rbf_and_grad


def part_major_perm(H, nh):
    """row index p*H + head*hd + c  <-  reference row head*3*hd + p*hd + c   (torchmd_et.py:275,327)"""
    hd = H // nh
    idx = torch.arange(3 * H).view(nh, 3, hd)           # [head, part, c] -> reference row
    return idx.permute(1, 0, 2).reshape(-1)


def pack_et(sd, hp, pre="representation_model."):
    H, nh, L = hp["embedding_dimension"], hp["num_heads"], hp["num_layers"]
    perm = part_major_perm(H, nh)
    W = dict(emb=sd[pre + "embedding.weight"], layers=[])
    if hp["neighbor_embedding"]:
        ne = pre + "neighbor_embedding."
        W["ne"] = dict(emb=sd[ne + "embedding.weight"], dp=(sd[ne + "distance_proj.weight"], sd[ne + "distance_proj.bias"]),
                       comb=(sd[ne + "combine.weight"], sd[ne + "combine.bias"]))
    di = hp["distance_influence"]
    for l in range(L):
        lp = pre + "attention_layers.%d." % l
        Lw = dict(ln=(sd[lp + "layernorm.weight"], sd[lp + "layernorm.bias"]),
                  # one [5H,H] projection: q | k | v(part-major)
                  qkv=(torch.cat([sd[lp + "q_proj.weight"], sd[lp + "k_proj.weight"], sd[lp + "v_proj.weight"][perm]]),
                       torch.cat([sd[lp + "q_proj.bias"], sd[lp + "k_proj.bias"], sd[lp + "v_proj.bias"][perm]])),
                  vp=sd[lp + "vec_proj.weight"],
                  o=(sd[lp + "o_proj.weight"], sd[lp + "o_proj.bias"]),
                  dk=(sd[lp + "dk_proj.weight"], sd[lp + "dk_proj.bias"]) if di in ("keys", "both") else None,
                  dv=(sd[lp + "dv_proj.weight"][perm], sd[lp + "dv_proj.bias"][perm]) if di in ("values", "both") else None)
        W["layers"].append(Lw)
    W["out_ln"] = (sd[pre + "out_norm.weight"], sd[pre + "out_norm.bias"])
    W["head"] = []
    for b in (0, 1):
        hp_ = "output_model.output_network.%d." % b
        W["head"].append(dict(w1=sd[hp_ + "vec1_proj.weight"], w2=sd[hp_ + "vec2_proj.weight"],
                              u0=(sd[hp_ + "update_net.0.weight"], sd[hp_ + "update_net.0.bias"]),
                              u2=(sd[hp_ + "update_net.2.weight"], sd[hp_ + "update_net.2.bias"])))
    if hp["rbf_type"] == "expnorm":
        W["rbf"] = (sd[pre + "distance_expansion.means"], sd[pre + "distance_expansion.betas"])
    else:
        W["rbf"] = (sd[pre + "distance_expansion.offset"], sd[pre + "distance_expansion.coeff"])
    W["std"] = sd.get("std", torch.tensor(1.0))
    W["mean"] = sd.get("mean", torch.tensor(0.0))
    return W


def edge_filter(d, hp, rbfp, lin, act, n_out):
    """act(Lin(rbf(d))) and d/dd (NO cutoff factor: torchmd_et.py:282-291); ones/zeros when the filter is absent."""
    if lin is None:
        return torch.ones(d.shape[0], n_out, dtype=d.dtype), torch.zeros(d.shape[0], n_out, dtype=d.dtype)
    r, dr = rbf_and_grad(d, hp, *rbfp)
    h, t = F.linear(r, *lin), F.linear(dr, lin[0])
    a, g = act_fn(act, h)
    return a, t * g


def head_sum(x, nh):
    """[...,H] -> per-head sums broadcast back to [...,H]"""
    sh = x.shape
    s = x.reshape(sh[:-1] + (nh, sh[-1] // nh)).sum(-1, keepdim=True)
    return s.expand(sh[:-1] + (nh, sh[-1] // nh)).reshape(sh)


# ------------------------------------------------------------------ gated equivariant block (models/utils.py:496-522)
def gated_fwd(Wb, x, v, act, out_ch, scalar_act):
    vb = F.linear(v, Wb["w1"])                                   # [N,3,Hin]
    n2 = (vb * vb).sum(1)
    mask = (vb != 0).reshape(vb.shape[0], -1).any(1)
    vec1 = torch.where(mask[:, None], torch.sqrt(n2), torch.zeros_like(n2))
    vec2 = F.linear(v, Wb["w2"])                                 # [N,3,out]
    h0 = F.linear(torch.cat([x, vec1], -1), *Wb["u0"])
    a0, g0 = act_fn(act, h0)
    h2 = F.linear(a0, *Wb["u2"])
    xo, gate = h2[:, :out_ch], h2[:, out_ch:]
    vo = gate[:, None, :] * vec2
    gx = torch.ones_like(xo)
    if scalar_act:
        xo, gx = act_fn(act, xo)
    return xo, vo, dict(vb=vb, vec1=vec1, vec2=vec2, g0=g0, gate=gate, gx=gx, hin=x.shape[1])


def gated_bwd(Wb, S, g_xo, g_vo):
    g_h2 = torch.cat([g_xo * S["gx"], (g_vo * S["vec2"]).sum(1)], -1)
    g_vec2 = g_vo * S["gate"][:, None, :]
    g_h0 = (g_h2 @ Wb["u2"][0]) * S["g0"]
    g_cat = g_h0 @ Wb["u0"][0]
    g_x, g_vec1 = g_cat[:, :S["hin"]], g_cat[:, S["hin"]:]
    inv = torch.where(S["vec1"] > 0, 1 / torch.where(S["vec1"] > 0, S["vec1"], torch.ones_like(S["vec1"])),
                      torch.zeros_like(S["vec1"]))
    g_vb = (g_vec1 * inv)[:, None, :] * S["vb"]
    g_v = g_vb @ Wb["w1"] + g_vec2 @ Wb["w2"]
    return g_x, g_v


# ------------------------------------------------------------------ the staged pipeline
def energy_forces_staged(sd, hp, z, pos, batch, ei, shift, n_mol, keep=None):
    """ei/shift as returned by oracle.model.build_edges (ei[0] = source j, ei[1] = target i for ET).
    Returns (y[M,1], forces[N,3])."""
    keep = {} if keep is None else keep
    W = pack_et(sd, hp)
    act, attn_act = hp["activation"], hp["attn_activation"]
    lo, hi = hp["cutoff_lower"], hp["cutoff_upper"]
    H, nh, n = hp["embedding_dimension"], hp["num_heads"], z.shape[0]
    row_ptr, tgt, col, perm = csr_from_edges(torch.stack([ei[1], ei[0]]), n)          # rows = targets
    dvec = pos[tgt] - pos[col] - shift[perm]                                            # pos_i - pos_j (min image)
    selfm = tgt == col
    d = torch.where(selfm, torch.zeros_like(dvec[:, 0]), dvec.norm(dim=1))
    u = torch.where(selfm[:, None], torch.zeros_like(dvec), dvec / torch.where(selfm, torch.ones_like(d), d)[:, None])
    dh = -u                                                                             # reference d_ij
    C, dC = cutoff_and_grad(d, lo, hi)
    keep.update(row_ptr=row_ptr, col=col, u=u, d=d)
    zero_x = lambda: torch.zeros(n, H, dtype=pos.dtype)                                 # noqa: E731

    # ---------------- forward
    x = W["emb"][z]
    if "ne" in W:
        r, dr = rbf_and_grad(d, hp, *W["rbf"])
        lin = F.linear(r, *W["ne"]["dp"])
        fne = lin * C[:, None]                                                          # models/utils.py:98-99
        dfne = F.linear(dr, W["ne"]["dp"][0]) * C[:, None] + lin * dC[:, None]
        ns = (~selfm).to(pos.dtype)[:, None]                                            # self loops removed, :92-96
        xn = zero_x().index_add(0, tgt, W["ne"]["emb"][z[col]] * fne * ns)
        x = F.linear(torch.cat([x, xn], 1), *W["ne"]["comb"])
        keep.update(xn=xn)
    vec = torch.zeros(n, 3, H, dtype=pos.dtype)
    keep.update(x0=x)
    saved = []
    for l, L in enumerate(W["layers"]):
        xl, ln = layer_norm_fwd(x, *L["ln"])
        qkv = F.linear(xl, *L["qkv"])
        q, k, v = qkv[:, :H], qkv[:, H:2 * H], qkv[:, 2 * H:].reshape(n, 3, H)
        vp = F.linear(vec, L["vp"])                                                     # [N,3,3H]
        vec1, vec2, vec3 = vp[..., :H], vp[..., H:2 * H], vp[..., 2 * H:]
        vec_dot = (vec1 * vec2).sum(1)
        dk, ddk = edge_filter(d, hp, W["rbf"], L["dk"], act, H)
        dv, ddv = edge_filter(d, hp, W["rbf"], L["dv"], act, 3 * H)
        dv, ddv = dv.view(-1, 3, H), ddv.view(-1, 3, H)
        s = head_sum(q[tgt] * k[col] * dk, nh)                                          # [E,H] (constant within a head)
        a, ga = act_fn(attn_act, s)
        attn = a * C[:, None]
        vj = v[col] * dv                                                                # [E,3,H]
        xmsg = vj[:, 0] * attn
        vmsg = vec[col] * vj[:, 1, None, :] + vj[:, 2, None, :] * dh[:, :, None]
        xa = zero_x().index_add(0, tgt, xmsg)
        va = torch.zeros_like(vec).index_add(0, tgt, vmsg)
        o = F.linear(xa, *L["o"])
        o1, o2, o3 = o[:, :H], o[:, H:2 * H], o[:, 2 * H:]
        dx = vec_dot * o2 + o3
        dvc = vec3 * o1[:, None, :] + va
        saved.append(dict(x=x, ln=ln, q=q, k=k, v=v, vec=vec, vp=vp, o=o, dk=dk, ddk=ddk, dv=dv, ddv=ddv))
        keep.update({"q%d" % l: q, "k%d" % l: k, "v%d" % l: v, "xa%d" % l: xa, "va%d" % l: va})
        x, vec = x + dx, vec + dvc
        keep.update({"x%d" % (l + 1): x, "vec%d" % (l + 1): vec})
    xo, lno = layer_norm_fwd(x, *W["out_ln"])
    h1, v1, S1 = gated_fwd(W["head"][0], xo, vec, act, H // 2, True)
    h2, v2, S2 = gated_fwd(W["head"][1], h1, v1, act, 1, False)
    e_atom = h2 * W["std"]
    y = torch.zeros(n_mol, 1, dtype=pos.dtype).index_add(0, batch, e_atom) + W["mean"]
    keep.update(e_atom=e_atom, xo=xo)

    # ---------------- backward (dL/dy = 1)
    g_h1, g_v1 = gated_bwd(W["head"][1], S2, torch.ones(n, 1, dtype=pos.dtype) * W["std"], torch.zeros_like(v2))
    g_xo, g_vec = gated_bwd(W["head"][0], S1, g_h1, g_v1)
    g_x = layer_norm_bwd(g_xo, W["out_ln"][0], lno)
    gpos = torch.zeros(n, 3, dtype=pos.dtype)
    inv_d = torch.where(selfm, torch.zeros_like(d), 1 / torch.where(selfm, torch.ones_like(d), d))

    def pos_grad(g_d, g_dh, dhat):
        """dL/d(vec_e), vec_e = d * dhat:  g_d dhat + (g_dh - (g_dh.dhat) dhat)/d"""
        return g_d[:, None] * dhat + (g_dh - (g_dh * dhat).sum(1, keepdim=True) * dhat) * inv_d[:, None]

    for l in reversed(range(len(W["layers"]))):
        L, S = W["layers"][l], saved[l]
        q, k, v, vec_l, vp, o = S["q"], S["k"], S["v"], S["vec"], S["vp"], S["o"]
        dk, ddk, dv, ddv = S["dk"], S["ddk"], S["dv"], S["ddv"]
        vec1, vec2, vec3 = vp[..., :H], vp[..., H:2 * H], vp[..., 2 * H:]
        o1, o2 = o[:, :H], o[:, H:2 * H]
        # node stage: dx = vec_dot*o2 + o3 ; dvec = vec3*o1 + va
        vec_dot = (vec1 * vec2).sum(1)
        g_o = torch.cat([(g_vec * vec3).sum(1), g_x * vec_dot, g_x], 1)                 # [N,3H]  (o1 | o2 | o3)
        g_vdot = g_x * o2
        g_vp = torch.cat([g_vdot[:, None, :] * vec2, g_vdot[:, None, :] * vec1, g_vec * o1[:, None, :]], -1)
        g_va = g_vec                                                                    # [N,3,H]
        g_xa = g_o @ L["o"][0]                                                          # [N,H]
        g_vec_in = g_vec + g_vp @ L["vp"]                                               # residual + vec_proj^T
        # edge stage, direction A: entry (i <- j) seen from its target i  -> q_i adjoint, position adjoints
        s = head_sum(q[tgt] * k[col] * dk, nh)
        a, ga = act_fn(attn_act, s)
        attn = a * C[:, None]
        vj = v[col] * dv
        g_attn = head_sum(g_xa[tgt] * vj[:, 0], nh)                                     # per head
        g_s = g_attn * ga * C[:, None]
        g_q = zero_x().index_add(0, tgt, g_s * k[col] * dk)
        g_v1 = (g_va[tgt] * vec_l[col]).sum(1)                                          # [E,H]
        g_v2 = (g_va[tgt] * dh[:, :, None]).sum(1)
        hd = H // nh
        gdA = (g_s * q[tgt] * k[col] * ddk + g_xa[tgt] * attn * v[col][:, 0] * ddv[:, 0]
               + g_v1 * v[col][:, 1] * ddv[:, 1] + g_v2 * v[col][:, 2] * ddv[:, 2]
               + g_attn * a / hd * dC[:, None]).sum(1)
        gdhA = (g_va[tgt] * vj[:, 2, None, :]).sum(2)                                   # [E,3]
        # direction B: the reverse edge (j <- i) of the same entry, seen from its source i -> k_i, v_i, vec_i adjoints
        dhB = -dh
        sB = head_sum(q[col] * k[tgt] * dk, nh)
        aB, gaB = act_fn(attn_act, sB)
        attnB = aB * C[:, None]
        viB = v[tgt] * dv                                                               # v of the source (= i) * dv
        g_attnB = head_sum(g_xa[col] * viB[:, 0], nh)
        g_sB = g_attnB * gaB * C[:, None]
        g_k = zero_x().index_add(0, tgt, g_sB * q[col] * dk)
        g_v1B = (g_va[col] * vec_l[tgt]).sum(1)
        g_v2B = (g_va[col] * dhB[:, :, None]).sum(1)
        g_v = torch.zeros_like(v).index_add(0, tgt, torch.stack([g_xa[col] * attnB, g_v1B, g_v2B], 1) * dv)
        g_vec_in = g_vec_in.index_add(0, tgt, g_va[col] * viB[:, 1, None, :])
        gdB = (g_sB * q[col] * k[tgt] * ddk + g_xa[col] * attnB * v[tgt][:, 0] * ddv[:, 0]
               + g_v1B * v[tgt][:, 1] * ddv[:, 1] + g_v2B * v[tgt][:, 2] * ddv[:, 2]
               + g_attnB * aB / hd * dC[:, None]).sum(1)
        gdhB = (g_va[col] * viB[:, 2, None, :]).sum(2)
        # vec_e = pos_j - pos_i for A (atom i gets -G_A), vec_e' = pos_i - pos_j for B (atom i gets +G_B)
        gpos = gpos.index_add(0, tgt, -pos_grad(gdA, gdhA, dh) + pos_grad(gdB, gdhB, dhB))
        # projections
        g_qkv = torch.cat([g_q, g_k, g_v.reshape(n, 3 * H)], 1)
        g_xl = g_qkv @ L["qkv"][0]
        g_x = g_x + layer_norm_bwd(g_xl, L["ln"][0], S["ln"])
        g_vec = g_vec_in
    if "ne" in W:
        g_cat = g_x @ W["ne"]["comb"][0]
        g_xn = g_cat[:, H:]
        ns = (~selfm).to(pos.dtype)
        nemb = W["ne"]["emb"]
        gd = ((g_xn[tgt] * nemb[z[col]] + g_xn[col] * nemb[z[tgt]]) * dfne).sum(1) * ns
        # both directions only depend on d: atom i receives -g_d(e) dh + g_d(e') dhB = -(g_dA + g_dB) dh
        gpos = gpos.index_add(0, tgt, -gd[:, None] * dh)
    return y, -gpos
