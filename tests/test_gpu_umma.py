"""Pins the tcgen05 operand-image conventions of torch-ngp_b200/csrc/umma.cuh on real hardware.

The shared-memory byte images are built HERE in numpy (128B-swizzled tiles), the diagnostics entry
ngp_umma_selftest just issues the described tcgen05.mma sequence and dumps TMEM; the result must
equal a plain fp32 matmul of the fp16 operands.  Covers the K-major view (forward products) and the
MN-major view (transposed products of the backward pass), single and double 64-wide slabs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def swz_image(mat):
    """[R, C<=64] fp16 -> R*128 bytes: 16-byte chunk c of row r at chunk position c ^ (r & 7)."""
    R, C = mat.shape
    assert C % 8 == 0 and C <= 64
    img = np.zeros((R, 8, 8), np.float16)  # row, chunk position, 8 halves
    for c in range(C // 8):
        pos = (c ^ (np.arange(R) & 7))
        img[np.arange(R), pos] = mat[:, c * 8:(c + 1) * 8]
    return img.reshape(-1).view(np.uint8)


def idesc(M, N, a_mn, b_mn):
    return (1 << 4) | (int(a_mn) << 15) | (int(b_mn) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def run(a_img, b_img, N, nk, desc, a=(0, 1024, 32), b=(0, 1024, 32)):
    import ngp_abi as abi
    ta, tb = torch.from_numpy(a_img.copy()).to(DEV), torch.from_numpy(b_img.copy()).to(DEV)
    D = torch.full((128, N), float("nan"), device=DEV)
    abi.check(abi.lib.ngp_umma_selftest(ta.data_ptr(), ta.numel(), tb.data_ptr(), tb.numel(), D.data_ptr(), N, nk, desc,
                                        a[0], a[1], a[2], b[0], b[1], b[2], abi.stream()))
    torch.cuda.synchronize()
    return D.cpu().numpy()


@pytest.mark.parametrize("N,K", [(16, 16), (64, 32), (64, 64), (128, 64), (256, 48), (32, 64)])
def test_kmajor_a_kmajor_b(N, K):
    rng = np.random.default_rng(N * 100 + K)
    A = rng.standard_normal((128, K)).astype(np.float16)
    B = rng.standard_normal((N, K)).astype(np.float16)
    D = run(swz_image(A), swz_image(B), N, K // 16, idesc(128, N, False, False))
    ref = A.astype(np.float32) @ B.astype(np.float32).T
    assert np.allclose(D, ref, rtol=1e-3, atol=1e-3), np.abs(D - ref).max()


@pytest.mark.parametrize("N,K", [(16, 16), (32, 64), (64, 16), (64, 64), (64, 128)])
def test_kmajor_a_mnmajor_b(N, K):
    """B given as Bm [K rows, N cols] (e.g. a weight matrix [out=K, in=N] used for G @ W)."""
    rng = np.random.default_rng(N * 100 + K + 7)
    Kp = min(K, 64)
    A = rng.standard_normal((128, K)).astype(np.float16)
    Bm = rng.standard_normal((K, N)).astype(np.float16)
    ref = A.astype(np.float32) @ Bm.astype(np.float32)
    if K <= 64:
        D = run(swz_image(A), swz_image(Bm), N, K // 16, idesc(128, N, False, True), b=(0, 1024, 2048))
        assert np.allclose(D, ref, rtol=1e-3, atol=1e-3), np.abs(D - ref).max()
    else:  # two K-major A tiles: first 64 then next 64 -- two launches' worth, checked by linearity
        D0 = run(swz_image(A[:, :64]), swz_image(Bm[:64]), N, 4, idesc(128, N, False, True), b=(0, 1024, 2048))
        D1 = run(swz_image(A[:, 64:]), swz_image(Bm[64:]), N, 4, idesc(128, N, False, True), b=(0, 1024, 2048))
        assert np.allclose(D0 + D1, ref, rtol=1e-3, atol=1e-3)
    del Kp


def test_kmajor_a_mnmajor_b_two_slabs():
    """N = 128 as two 64-wide MN slabs, LBO = slab stride."""
    rng = np.random.default_rng(5)
    K = 32
    A = rng.standard_normal((128, K)).astype(np.float16)
    Bm = rng.standard_normal((K, 128)).astype(np.float16)
    img = np.concatenate([swz_image(Bm[:, :64]), swz_image(Bm[:, 64:])])
    D = run(swz_image(A), img, 128, K // 16, idesc(128, 128, False, True), b=(K * 128, 1024, 2048))
    ref = A.astype(np.float32) @ Bm.astype(np.float32)
    assert np.allclose(D, ref, rtol=1e-3, atol=1e-3), np.abs(D - ref).max()


@pytest.mark.parametrize("M,N,K", [(64, 64, 128), (64, 32, 128), (16, 64, 128), (64, 64, 16), (128, 64, 128)])
def test_mnmajor_a_mnmajor_b(M, N, K):
    """Weight-gradient product D = Gm^T @ Hm with Gm [K rows, M], Hm [K rows, N] (K = batch rows of the tile).
    M < 128: the second 64-wide slab of A aliases the first (LBO = 0), rows >= 64 of D are ignored."""
    rng = np.random.default_rng(M + N + K)
    Gm = rng.standard_normal((K, M)).astype(np.float16)
    Hm = rng.standard_normal((K, N)).astype(np.float16)
    ref = Gm.astype(np.float32).T @ Hm.astype(np.float32)
    if M <= 64:
        Gpad = np.zeros((K, 64), np.float16)
        Gpad[:, :M] = Gm
        a_img, a_lbo = swz_image(Gpad), 0
    else:
        a_img, a_lbo = np.concatenate([swz_image(Gm[:, :64]), swz_image(Gm[:, 64:])]), K * 128
    D = run(a_img, swz_image(Hm), N, K // 16, idesc(128, N, True, True), a=(a_lbo, 1024, 2048), b=(0, 1024, 2048))
    assert np.allclose(D[:M], ref, rtol=1e-3, atol=2e-3), np.abs(D[:M] - ref).max()
    if M <= 64:
        assert np.allclose(D[64:64 + M], ref, rtol=1e-3, atol=2e-3)  # aliased slab
