"""CPU tests of the compressed block-tridiagonal layout helpers and the multi-GPU partition."""
import numpy as np
import pytest

from gbd_pcg_b200 import layout, partition, synth


@pytest.mark.parametrize("nx,N", [(1, 2), (2, 3), (5, 4), (14, 32)])
def test_pack_unpack_roundtrip(nx, N):
    rng = np.random.default_rng(0)
    n = nx * N
    A = np.zeros((n, n))
    for k in range(N):
        for kc in (k - 1, k, k + 1):
            if 0 <= kc < N:
                A[k * nx:(k + 1) * nx, kc * nx:(kc + 1) * nx] = rng.standard_normal((nx, nx))
    M = layout.pack_dense(A, nx, N)
    assert M.size == layout.mat_elems(nx, N)
    assert np.array_equal(layout.unpack_dense(M, nx, N), A)
    # padding blocks are where the layout says (left of row 0, right of row N-1)
    assert not M[:nx * nx].any() and not M[-nx * nx:].any()
    # element addressing of reference utils.cuh:58-84: M[k*3nx^2 + j*nx^2 + c*nx + r] = A[k*nx+r, (k+j-1)*nx+c]
    k, j, c, r = N - 1, 0, nx - 1, 0
    assert M[k * 3 * nx * nx + j * nx * nx + c * nx + r] == A[k * nx + r, (k + j - 1) * nx + c]


def test_pack_blocks_equals_pack_dense():
    diag, lower, _ = synth.lqr_blocks(6, 3, 7, 1)
    n = 42
    A = np.zeros((n, n))
    for k in range(7):
        A[k * 6:(k + 1) * 6, k * 6:(k + 1) * 6] = diag[k]
        if k < 6:
            A[(k + 1) * 6:(k + 2) * 6, k * 6:(k + 1) * 6] = lower[k]
            A[k * 6:(k + 1) * 6, (k + 1) * 6:(k + 2) * 6] = lower[k].T
    assert np.array_equal(layout.pack_blocks(diag, lower), layout.pack_dense(A, 6, 7))


def test_identity_and_bytes():
    I = layout.identity_compressed(3, 4, np.float64)
    assert np.array_equal(layout.unpack_dense(I, 3, 4), np.eye(12))
    # algorithmic bytes per solve
    assert layout.algorithmic_bytes(2, 3, 4) == 296
    assert layout.algorithmic_bytes(14, 32, 4) == 152768
    assert layout.algorithmic_bytes(14, 64, 8) == 617344
    assert layout.algorithmic_bytes(12, 512, 4) == 1840896


def test_generator_is_deterministic_and_seeded():
    a = synth.lqr_schur(14, 7, 32, 5)
    b = synth.lqr_schur(14, 7, 32, 5)
    c = synth.lqr_schur(14, 7, 32, 6)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert not np.array_equal(a[0], c[0])


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("batch", [0, 1, 7, 8192, 8191])
def test_partition_is_contiguous_and_complete(world, batch):
    slabs = [partition.slab(r, world, batch) for r in range(world)]
    assert slabs[0][0] == 0 and slabs[-1][1] == batch
    assert all(slabs[i][1] == slabs[i + 1][0] for i in range(world - 1))
    sizes = [hi - lo for lo, hi in slabs]
    assert max(sizes) - min(sizes) <= 1
    for i in range(0, batch, max(1, batch // 50)):
        r = partition.owner(i, world, batch)
        assert slabs[r][0] <= i < slabs[r][1]
