"""The arithmetic of the 8 x 8 diagonal block in k_solve_cam (csrc/kernels.cuh, "(1) diagonal block"), restated in numpy.

The kernel eliminates the block fraction-free -- a_ic <- (a_ic p_j - a_ij a_cj) 2^-e(p_j) -- so that no division or
reciprocal square root sits on the chain of dependent pivots, and recovers the Cholesky factor afterwards from
L_ij = a_ij^(j) / sqrt(p_j q_j), 1 / L_jj = q_j / sqrt(p_j q_j), q_{j+1} = q_j p_j 2^-e(p_j).  These tests pin the algebra
(the GPU parity tests pin the kernel): the result is a Cholesky factor to working precision on blocks of any scale and
conditioning the solver meets, the running product q stays within [1, 256], and the guard that sends a block to the textbook
elimination fires exactly for non-positive pivots."""
import numpy as np


def fraction_free_cholesky(A):
    n = len(A)
    a = np.array(A, dtype=np.float64)
    p, s = np.ones(n), np.ones(n)
    for j in range(n):
        p[j] = a[j, j]
        _, e = np.frexp(p[j])
        s[j] = 2.0 ** (1 - e)                         # p s in [1, 2): what the exponent arithmetic of the kernel produces
        col = a[:, j].copy()
        for c in range(j + 1, n):
            a[:, c] = (a[:, c] * p[j] - col * col[c]) * s[j]
    q, ok = 1.0, True
    L, linv, qs = np.zeros_like(a), np.zeros(n), []
    for j in range(n):
        pq = p[j] * q
        ok = ok and p[j] > 0.0 and 1e-250 < pq < 1e250
        qs.append(q)
        t = 1.0 / np.sqrt(pq) if pq > 0 else np.nan
        L[:, j] = a[:, j] * t
        linv[j] = q * t
        q = pq * s[j]
    return np.tril(L), linv, ok, qs


def test_matches_cholesky_over_scales_and_condition_numbers():
    rng = np.random.default_rng(3)
    worst = 0.0
    for trial in range(3000):
        n = 8
        B = rng.standard_normal((n, n + 2)) * 10.0 ** rng.uniform(-6, 4, (n, 1))
        A = B @ B.T + np.eye(n) * 10.0 ** rng.uniform(-14, -2)
        if trial % 2:                                  # the solver's systems are Jacobi-scaled: unit-ish diagonal
            d = 1.0 / np.sqrt(np.diag(A))
            A = A * d[:, None] * d[None, :]
        try:
            L0 = np.linalg.cholesky(A)
        except np.linalg.LinAlgError:
            continue
        L, linv, ok, qs = fraction_free_cholesky(A)
        assert ok
        assert max(qs) <= 256.0 and min(qs) >= 1.0
        back = np.abs(L @ L.T - A).max() / np.abs(A).max()
        back0 = np.abs(L0 @ L0.T - A).max() / np.abs(A).max()
        worst = max(worst, back)
        assert back <= 4e-15 and back <= 10 * max(back0, 2.3e-16)
        assert np.abs(linv * np.diag(L) - 1.0).max() <= 1e-15
    assert worst > 0.0


def test_guard_fires_for_indefinite_blocks_only():
    rng = np.random.default_rng(5)
    B = rng.standard_normal((8, 8))
    A = B @ B.T + 0.1 * np.eye(8)
    assert fraction_free_cholesky(A)[2]
    A2 = A.copy()
    A2[5, 5] -= 2.0 * np.linalg.eigvalsh(A).max()      # not positive definite any more
    assert not fraction_free_cholesky(A2)[2]
    assert not fraction_free_cholesky(A * 1e-280)[2]        # out of the product range: textbook elimination takes over
