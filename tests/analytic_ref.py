"""numpy statement of what MB2_TANGENT_ANALYTIC assembles (mimi_b200/csrc/mb2_fast.cu): the closed-form tangent
K_e[(i,j),(l,m)] = sum_q sum_{k,n} dN(q;i,k) (w detJ)_q dP_jk/dF_mn dN(q;l,n) of CompressibleOgdenNeoHookean
(reference materials.cpp:96-118 + 60-71: P = mu F + (-mu + lambda J (J - 1)) F^-T) and StVenantKirchhoff
(materials.cpp:73-94: P = F S). Test infrastructure: float64 numpy, no claim on operation order."""
import numpy as np


def pk1_and_tangent(mat, lam, mu, F):
    """P [d, d] and C[j, k, m, n] = dP_jk / dF_mn."""
    d = F.shape[0]
    I = np.eye(d)
    if mat == "stvk":
        E = 0.5 * (F.T @ F - I)
        S = lam * np.trace(E) * I + 2.0 * mu * E
        P = F @ S
        B = F @ F.T
        C = (np.einsum("jm,nk->jkmn", I, S) + lam * np.einsum("jk,mn->jkmn", F, F)
             + mu * (np.einsum("jn,mk->jkmn", F, F) + np.einsum("jm,kn->jkmn", B, I)))
        return P, C
    J = np.linalg.det(F)
    G = np.linalg.inv(F).T
    g = -mu + lam * J * (J - 1.0)
    P = mu * F + g * G
    C = (mu * np.einsum("jm,kn->jkmn", I, I) + lam * (2.0 * J - 1.0) * J * np.einsum("jk,mn->jkmn", G, G)
         - g * np.einsum("jn,mk->jkmn", G, G))
    return P, C


def element(mat, lam, mu, dim, nd, ue, dN_dX, wdetJ):
    """ue [dim*nd] component-major; dN_dX [nq, dim*nd] = QuadData::dN_dX column-major (k*nd + i); -> R [tdof], K [tdof, tdof]
    with row = i + j nd, column = l + m nd."""
    nq = wdetJ.shape[0]
    U = ue.reshape(dim, nd)                       # U[c, a]
    dN = dN_dX.reshape(nq, dim, nd)               # dN[q, k, i]
    R = np.zeros((dim, nd))
    K = np.zeros((dim, nd, dim, nd))              # K[j, i, m, l]
    for q in range(nq):
        F = np.eye(dim) + U @ dN[q].T             # F[c, k] = delta + sum_a U[c, a] dN[k, a]
        P, C = pk1_and_tangent(mat, lam, mu, F)
        R += wdetJ[q] * np.einsum("ki,jk->ji", dN[q], P)
        K += wdetJ[q] * np.einsum("ki,jkmn,nl->jiml", dN[q], C, dN[q])
    return R.reshape(-1), K.reshape(dim * nd, dim * nd)


def assemble(mat, lam, mu, D, u, gf, A0):
    """r = sum R_e, A = A0 + gf * sum K_e through the reference's A_ids (column-major K_e: A_ids[c * tdof + r])."""
    dim, nd, nq, NE, tdof = D["dim"], D["nd"], D["nq"], D["NE"], D["tdof"]
    r = np.zeros(D["NV"])
    A = A0.copy()
    dN = D["dN_dX"].reshape(NE, nq, tdof)
    wd = (D["w"] * D["detJ"]).reshape(NE, nq)
    vd = D["v_dofs"].reshape(NE, tdof)
    ids = D["A_ids"].reshape(NE, tdof, tdof)      # [col][row]
    for e in range(NE):
        Re, Ke = element(mat, lam, mu, dim, nd, u[vd[e]], dN[e], wd[e])
        np.add.at(r, vd[e], Re)
        np.add.at(A, ids[e].T.reshape(-1), gf * Ke.reshape(-1))
    return r, A
