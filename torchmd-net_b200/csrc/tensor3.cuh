// tensor3.cuh -- per-(atom,channel) rank-2 tensor algebra in irreducible components.
//
// A channel of a TensorNet node feature is a 3x3 matrix X = I + A + S (tensornet.py:48-54).  It is stored as
// nine numbers  x = (t | a1 a2 a3 | s1 s2 s3 s4 s5):
//     I = t*1,   A = [[0,-a3,a2],[a3,0,-a1],[-a2,a1,0]],   S = [[s1,s3,s4],[s3,s2,s5],[s4,s5,-s1-s2]]
// so decompose_tensor() is free and the per-component channel mixes are plain GEMMs over rows.
// The same formulas, in torch, are checked against autograd in oracle/staged_tensornet.py.
#pragma once

#define T3_HD __host__ __device__ __forceinline__

T3_HD void t3_to_mat(const float x[9], float m[9]) {
    m[0] = x[0] + x[4]; m[1] = x[6] - x[3]; m[2] = x[7] + x[2];
    m[3] = x[6] + x[3]; m[4] = x[0] + x[5]; m[5] = x[8] - x[1];
    m[6] = x[7] - x[2]; m[7] = x[8] + x[1]; m[8] = x[0] - x[4] - x[5];
}
T3_HD void t3_from_mat(const float m[9], float x[9]) {
    const float t = (m[0] + m[4] + m[8]) * (1.0f / 3.0f);
    x[0] = t;
    x[1] = 0.5f * (m[7] - m[5]);
    x[2] = 0.5f * (m[2] - m[6]);
    x[3] = 0.5f * (m[3] - m[1]);
    x[4] = m[0] - t;
    x[5] = m[4] - t;
    x[6] = 0.5f * (m[1] + m[3]);
    x[7] = 0.5f * (m[2] + m[6]);
    x[8] = 0.5f * (m[5] + m[7]);
}
// adjoint of t3_to_mat: g = dL/dM -> dL/dx
T3_HD void t3_to_mat_T(const float g[9], float dx[9]) {
    dx[0] = g[0] + g[4] + g[8];
    dx[1] = g[7] - g[5];
    dx[2] = g[2] - g[6];
    dx[3] = g[3] - g[1];
    dx[4] = g[0] - g[8];
    dx[5] = g[4] - g[8];
    dx[6] = g[1] + g[3];
    dx[7] = g[2] + g[6];
    dx[8] = g[5] + g[7];
}
// adjoint of t3_from_mat: dL/dx -> dL/dM
T3_HD void t3_from_mat_T(const float dx[9], float g[9]) {
    const float k = (dx[0] - dx[4] - dx[5]) * (1.0f / 3.0f);
    g[0] = k + dx[4]; g[1] = 0.5f * (dx[6] - dx[3]); g[2] = 0.5f * (dx[7] + dx[2]);
    g[3] = 0.5f * (dx[6] + dx[3]); g[4] = k + dx[5]; g[5] = 0.5f * (dx[8] - dx[1]);
    g[6] = 0.5f * (dx[7] - dx[2]); g[7] = 0.5f * (dx[8] + dx[1]); g[8] = k;
}
// squared Frobenius norm of the three parts (tensor_norm, tensornet.py:66-67)
T3_HD void t3_qparts(const float x[9], float& qi, float& qa, float& qs) {
    qi = 3.0f * x[0] * x[0];
    qa = 2.0f * (x[1] * x[1] + x[2] * x[2] + x[3] * x[3]);
    const float s12 = x[4] + x[5];
    qs = x[4] * x[4] + x[5] * x[5] + s12 * s12 + 2.0f * (x[6] * x[6] + x[7] * x[7] + x[8] * x[8]);
}
T3_HD float t3_q(const float x[9]) {
    float a, b, c;
    t3_qparts(x, a, b, c);
    return a + b + c;
}
T3_HD void t3_qgrad(const float x[9], float g[9]) {
    g[0] = 6.0f * x[0];
    g[1] = 4.0f * x[1]; g[2] = 4.0f * x[2]; g[3] = 4.0f * x[3];
    g[4] = 4.0f * x[4] + 2.0f * x[5];
    g[5] = 4.0f * x[5] + 2.0f * x[4];
    g[6] = 4.0f * x[6]; g[7] = 4.0f * x[7]; g[8] = 4.0f * x[8];
}
// c = a*b (3x3 row-major);  c = a^T*b;  c = a*b^T
T3_HD void t3_mm(const float a[9], const float b[9], float c[9]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) c[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
}
T3_HD void t3_mtm(const float a[9], const float b[9], float c[9]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) c[3 * i + j] = a[i] * b[j] + a[3 + i] * b[3 + j] + a[6 + i] * b[6 + j];
}
T3_HD void t3_mmt(const float a[9], const float b[9], float c[9]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            c[3 * i + j] = a[3 * i] * b[3 * j] + a[3 * i + 1] * b[3 * j + 1] + a[3 * i + 2] * b[3 * j + 2];
}

// forward of  T = M*Y + Y*M (O(3)) or 2*Y*M (SO(3)) in irreps:  tau = from_mat(T)
T3_HD void t3_product(const float m[9], const float y[9], bool so3, float tau[9]) {
    float Mm[9], Ym[9], A[9], B[9];
    t3_to_mat(m, Mm);
    t3_to_mat(y, Ym);
    t3_mm(Ym, Mm, B);
    if (so3) {
#pragma unroll
        for (int k = 0; k < 9; ++k) A[k] = 2.0f * B[k];
    } else {
        t3_mm(Mm, Ym, A);
#pragma unroll
        for (int k = 0; k < 9; ++k) A[k] += B[k];
    }
    t3_from_mat(A, tau);
}
// adjoint: given dtau, returns dm, dy
T3_HD void t3_product_bwd(const float m[9], const float y[9], bool so3, const float dtau[9], float dm[9], float dy[9]) {
    float Mm[9], Ym[9], G[9], A[9], B[9];
    t3_to_mat(m, Mm);
    t3_to_mat(y, Ym);
    t3_from_mat_T(dtau, G);
    if (so3) {  // T = 2 Y M : dM = 2 Y^T G, dY = 2 G M^T
        t3_mtm(Ym, G, A);
        t3_mmt(G, Mm, B);
#pragma unroll
        for (int k = 0; k < 9; ++k) { A[k] *= 2.0f; B[k] *= 2.0f; }
    } else {    // T = M Y + Y M : dM = G Y^T + Y^T G, dY = M^T G + G M^T
        float C[9];
        t3_mmt(G, Ym, A);
        t3_mtm(Ym, G, C);
#pragma unroll
        for (int k = 0; k < 9; ++k) A[k] += C[k];
        t3_mtm(Mm, G, B);
        t3_mmt(G, Mm, C);
#pragma unroll
        for (int k = 0; k < 9; ++k) B[k] += C[k];
    }
    t3_to_mat_T(A, dm);
    t3_to_mat_T(B, dy);
}
// x/(q(x)+1) and its adjoint
T3_HD void t3_normalize(const float x[9], float xh[9]) {
    const float inv = 1.0f / (t3_q(x) + 1.0f);
#pragma unroll
    for (int k = 0; k < 9; ++k) xh[k] = x[k] * inv;
}
T3_HD void t3_normalize_bwd(const float x[9], const float dxh[9], float dx[9]) {
    const float inv = 1.0f / (t3_q(x) + 1.0f);
    float g[9], dot = 0.f;
    t3_qgrad(x, g);
#pragma unroll
    for (int k = 0; k < 9; ++k) dot += dxh[k] * x[k];
    const float s = dot * inv * inv;
#pragma unroll
    for (int k = 0; k < 9; ++k) dx[k] = dxh[k] * inv - s * g[k];
}
// sq = from_mat(to_mat(d)^2)  and adjoint contribution  dd += to_mat_T(Gs D^T + D^T Gs), Gs = from_mat_T(dxn)
T3_HD void t3_square(const float d[9], float sq[9]) {
    float D[9], D2[9];
    t3_to_mat(d, D);
    t3_mm(D, D, D2);
    t3_from_mat(D2, sq);
}
T3_HD void t3_square_bwd(const float d[9], const float dxn[9], float dd[9]) {
    float D[9], G[9], A[9], B[9];
    t3_to_mat(d, D);
    t3_from_mat_T(dxn, G);
    t3_mmt(G, D, A);
    t3_mtm(D, G, B);
#pragma unroll
    for (int k = 0; k < 9; ++k) A[k] += B[k];
    t3_to_mat_T(A, dd);
}
