"""The mathematics of the fused kernel's box-QP solver (csrc/gmr_stages.cuh::qp_solve_aba), restated in numpy and
checked against a dense solve on random kinematic trees (CPU test; the CUDA implementation itself is checked
against the oracle's QP in tests/test_gpu_parity.py).

mink's build_ik objective  1/2 x'Hx + c'x  with  H = d I + sum_t J_t'K_t J_t,  c = sum_t J_t'g_t,  J_t x = v_{link(t)},
v_i = v_parent(i) + S_i x_i, is the forward dynamics of a tree with link "inertias" K_t, bias forces g_t and joint
armature d, so the articulated-body recursion returns its stationary point without forming H:
  up:    X_i = [I_i | p_i] = own tasks + sum_children [I_c - U_c U_c' a_c | p_c - U_c b_c],
         U_c = I_c S_c,  a_c = 1 / (S_c'U_c + d),  b_c = a_c S_c'p_c      (locked at x_c = bound: a_c = 0, b_c = -bound)
  down:  x_c = -(U_c'v_parent) a_c - b_c,   v_c = v_parent + S_c x_c
  multiplier of a locked dof:  (Hx + c)_c = S_c' sum_{t below c} (K_t v + g_t) + d x_c
  floating base: its six dofs are solved as one block,  (S_b' I S_b + d) x_b = -S_b' p.
"""
import numpy as np
import pytest


def _random_tree(rng, n, base=0):
    """Dofs in DFS order; the first `base` dofs form a chain (a floating base) above everything else."""
    parent = [-1] + [i - 1 for i in range(1, base)]
    for i in range(max(base, 1), n):
        parent.append(int(rng.integers(max(base - 1, 0), i)))
    return parent[:n]


def _ancestors(parent, i):
    out = []
    while i >= 0:
        out.append(i)
        i = parent[i]
    return out


def _dense(parent, S, tasks, K, g, d):
    n = len(parent)
    H = d * np.eye(n)
    c = np.zeros(n)
    for t, Kt, gt in zip(tasks, K, g):
        a = _ancestors(parent, t)
        for k in a:
            c[k] += S[k] @ gt
            for l in a:
                H[k, l] += S[k] @ Kt @ S[l]
    return H, c


def _aba(parent, S, tasks, K, g, d, locked, bound, base=0):
    n = len(parent)
    I = [np.zeros((6, 6)) for _ in range(n)]
    p = [np.zeros(6) for _ in range(n)]
    for t, Kt, gt in zip(tasks, K, g):
        I[t] = I[t] + Kt
        p[t] = p[t] + gt
    U = [None] * n
    al = np.zeros(n)
    be = np.zeros(n)
    lowest = base - 1 if base else 0               # the base link gathers its X; the base dofs above it are not swept
    for i in range(n - 1, lowest - 1, -1):
        U[i] = I[i] @ S[i]
        if locked[i]:
            al[i], be[i] = 0.0, -bound[i]
        else:
            al[i] = 1.0 / (S[i] @ U[i] + d)
            be[i] = al[i] * (S[i] @ p[i])
        if parent[i] >= 0 and i > lowest:
            I[parent[i]] = I[parent[i]] + I[i] - np.outer(U[i], U[i]) * al[i]
            p[parent[i]] = p[parent[i]] + p[i] - U[i] * be[i]
    x = np.zeros(n)
    v = [np.zeros(6) for _ in range(n)]
    start = 0
    if base:
        Sb = np.stack([S[b] for b in range(base)], 1)                      # 6 x base
        M = Sb.T @ I[base - 1] @ Sb + d * np.eye(base)
        x[:base] = np.linalg.solve(M, -Sb.T @ p[base - 1])
        v[base - 1] = Sb @ x[:base]
        start = base
    for i in range(start, n):
        vp = v[parent[i]] if parent[i] >= 0 else np.zeros(6)
        x[i] = -(U[i] @ vp) * al[i] - be[i]
        v[i] = vp + S[i] * x[i]
    f = [Kt @ v[t] + gt for t, Kt, gt in zip(tasks, K, g)]
    lam = np.zeros(n)
    for i in range(n):
        F = sum((f[k] for k, t in enumerate(tasks) if i in _ancestors(parent, t)), np.zeros(6))
        lam[i] = S[i] @ F + d * x[i]
    return x, lam


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("base", [0, 6])
def test_articulated_body_sweeps_equal_the_dense_solve(seed, base):
    rng = np.random.default_rng(100 + seed)
    n = 24
    parent = _random_tree(rng, n, base)
    S = rng.normal(size=(n, 6))
    tasks = sorted(int(t) for t in rng.choice(np.arange(max(base - 1, 0), n), size=9, replace=False))
    # every leaf of the swept tree must carry a task (the kernel's schedule drops task-free subtrees): add them
    has_child = {p for p in parent if p >= 0}
    tasks = sorted(set(tasks) | {i for i in range(n) if i not in has_child})
    K, g = [], []
    for _ in tasks:
        A = rng.normal(size=(6, 6))
        K.append(A.T @ A)
        g.append(rng.normal(size=6))
    d = 0.5 + rng.random()
    locked = np.zeros(n, bool)
    locked[rng.choice(np.arange(base, n), size=4, replace=False)] = True       # base dofs are never limited
    bound = rng.normal(size=n) * 0.2
    H, c = _dense(parent, S, tasks, K, g, d)
    free = ~locked
    x_ref = np.where(locked, bound, 0.0)
    x_ref[free] = np.linalg.solve(H[np.ix_(free, free)], -c[free] - H[np.ix_(free, locked)] @ bound[locked])
    x, lam = _aba(parent, S, tasks, K, g, d, locked, bound, base)
    np.testing.assert_allclose(x, x_ref, atol=1e-11)
    np.testing.assert_allclose(lam, H @ x_ref + c, atol=1e-9)
    assert np.abs((H @ x_ref + c)[free]).max() < 1e-9


def test_two_tasks_on_one_dof_add_up():
    rng = np.random.default_rng(7)
    parent = [-1, 0, 1, 1]
    S = rng.normal(size=(4, 6))
    tasks = [2, 2, 3]
    K = [np.eye(6) * (k + 1) for k in range(3)]
    g = [rng.normal(size=6) for _ in range(3)]
    H, c = _dense(parent, S, tasks, K, g, 0.5)
    x, lam = _aba(parent, S, tasks, K, g, 0.5, np.zeros(4, bool), np.zeros(4))
    np.testing.assert_allclose(x, np.linalg.solve(H, -c), atol=1e-12)
    np.testing.assert_allclose(lam, 0.0, atol=1e-10)
