"""tcgen05 GEMM numerics vs a plain fp32 PyTorch reference of the same op (operands rounded to bf16)."""
import pytest
import torch

from consensusml_b200.ops import gemm as G

pytestmark = pytest.mark.gpu


def _ref_inputs(m, n, k, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(m, k, generator=g).cuda()
    b = torch.randn(n, k, generator=g).cuda()
    return a, b


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (128, 128, 512), (93, 93, 1984), (300, 200, 100), (257, 513, 72),
                                    (1000, 1984, 137), (2048, 2048, 1024)])
def test_scale_epilogue(m, n, k):
    a, b = _ref_inputs(m, n, k, 1)
    ab, _ = G.to_bf16_rows(a)
    bb, _ = G.to_bf16_rows(b)
    c = G.gemm_tn(ab, [bb], n, alpha=0.5)
    torch.cuda.synchronize()
    ref = 0.5 * (ab[:, :k].float() @ bb[:, :k].float().t())
    err = (c - ref).abs().max().item()
    assert err < 1e-3 * max(1.0, ref.abs().max().item()), err


def test_rbf_and_sqdist_epilogue():
    g = torch.Generator().manual_seed(2)
    x = (torch.randn(93, 1984, generator=g) * 2 + 3).cuda()
    z = (torch.randn(44, 1984, generator=g) * 2 + 3).cuda()
    xb, nx = G.to_bf16_rows(x)
    zb, nz = G.to_bf16_rows(z)
    xr, zr = xb[:, :1984].float(), zb[:, :1984].float()
    assert torch.allclose(nx, (xr * xr).sum(1), rtol=1e-5)
    d2 = G.gemm_tn(xb, [zb], 44, epilogue=G.EPI_SQDIST, norm_a=nx, norm_b=[nz])
    ref = torch.cdist(xr.double(), zr.double()) ** 2
    assert ((d2.double() - ref).abs() / ref.clamp_min(1)).max() < 1e-4
    gamma = 1.0 / 1984
    k = G.rbf_kernel(x, z, gamma)
    assert (k.double() - torch.exp(-gamma * ref)).abs().max() < 1e-4
    ks = G.rbf_kernel(x)
    assert (ks.diagonal() - 1).abs().max() < 2e-4   # |a|^2+|b|^2-2ab cancellation in fp32
    assert (ks - ks.t()).abs().max() < 1e-5


def test_multi_source_b_matches_single():
    a, b = _ref_inputs(200, 4 * 160, 256, 3)
    ab, na = G.to_bf16_rows(a)
    bb, nb = G.to_bf16_rows(b)
    whole = G.gemm_tn(ab, [bb], 640, epilogue=G.EPI_RBF, alpha=0.01, norm_a=na, norm_b=[nb])
    parts = [bb[i * 160:(i + 1) * 160].contiguous() for i in range(4)]
    norms = [nb[i * 160:(i + 1) * 160].contiguous() for i in range(4)]
    split = G.gemm_tn(ab, parts, 640, epilogue=G.EPI_RBF, alpha=0.01, norm_a=na, norm_b=norms)
    torch.cuda.synchronize()
    assert torch.equal(whole, split)
    # ragged last source
    split2 = G.gemm_tn(ab, parts, 600, epilogue=G.EPI_RBF, alpha=0.01, norm_a=na, norm_b=norms)
    assert torch.equal(whole[:, :600], split2)


@pytest.mark.parametrize("m,n,k", [(256, 256, 64), (512, 1024, 512), (300, 1000, 200), (1000, 1984, 137), (2304, 2560, 1000),
                                    (4096, 4096, 2048)])
def test_cta_pair_kernel_matches_reference(m, n, k):
    """cta_group::2 kernel (two SMs per 256 x 256 tile), forced for every shape including ragged edges."""
    a, b = _ref_inputs(m, n, k, 5)
    ab, na = G.to_bf16_rows(a)
    bb, nb = G.to_bf16_rows(b)
    ref = ab[:, :k].float() @ bb[:, :k].float().t()
    G.set_variant(2)
    try:
        c = G.gemm_tn(ab, [bb], n, alpha=0.5)
        d2 = G.gemm_tn(ab, [bb], n, epilogue=G.EPI_SQDIST, norm_a=na, norm_b=[nb])
        torch.cuda.synchronize()
    finally:
        G.set_variant(0)
    scale = max(1.0, ref.abs().max().item())
    assert (c - 0.5 * ref).abs().max().item() < 1e-3 * scale
    d2ref = (na[:, None] + nb[None, :] - 2 * ref).clamp_min(0)
    assert (d2 - d2ref).abs().max().item() < 2e-3 * scale
    G.set_variant(1)
    try:
        c1 = G.gemm_tn(ab, [bb], n, alpha=0.5)
        torch.cuda.synchronize()
    finally:
        G.set_variant(0)
    assert (c - c1).abs().max().item() < 1e-3 * scale


def test_cta_pair_multi_source():
    a, b = _ref_inputs(512, 4 * 384, 256, 6)
    ab, na = G.to_bf16_rows(a)
    bb, nb = G.to_bf16_rows(b)
    parts = [bb[i * 384:(i + 1) * 384].contiguous() for i in range(4)]
    norms = [nb[i * 384:(i + 1) * 384].contiguous() for i in range(4)]
    ref = torch.exp(-0.01 * (na[:, None] + nb[None, :] - 2 * (ab.float() @ bb.float().t())).clamp_min(0))
    G.set_variant(2)
    try:
        split = G.gemm_tn(ab, parts, 1500, epilogue=G.EPI_RBF, alpha=0.01, norm_a=na, norm_b=norms)
        torch.cuda.synchronize()
    finally:
        G.set_variant(0)
    assert (split - ref[:, :1500]).abs().max().item() < 1e-4


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("n_src,rows,k,rank", [(8, 750, 2500, 3), (4, 1024, 512, 0), (2, 300, 200, 1)])
def test_staged_allgather_gemm_with_local_sources(n_src, rows, k, rank, variant):
    """The fused all-gather -> GEMM with every 'peer' block living on this GPU: copier CTAs stage the blocks, GEMM CTAs
    (single-CTA kernel or CTA pairs) consume tiles as their flags arrive.  Same kernel path as the multi-GPU run."""
    g = torch.Generator().manual_seed(7)
    a = torch.randn(rows, k, generator=g).cuda()
    blocks = [torch.randn(rows, k, generator=g).cuda() for _ in range(n_src)]
    blocks[rank] = a
    ab, na = G.to_bf16_rows(a)
    bb = [G.to_bf16_rows(b) for b in blocks]
    ld = ab.shape[1]
    n_total = n_src * rows
    stage = torch.empty(n_total, ld, dtype=torch.bfloat16, device="cuda")
    snorm = torch.empty(n_total, device="cuda")
    from consensusml_b200 import _native
    flags = torch.zeros(_native.cuda().cml_gemm_staged_flags(n_src, rows), dtype=torch.int32, device="cuda")
    ref = torch.cat([ab.float() @ b.float().t() for b, _ in bb], 1)
    d2 = (na[:, None] + torch.cat([nb for _, nb in bb])[None, :] - 2 * ref).clamp_min(0)
    G.set_variant(variant)
    try:
        for epoch in (1, 2):
            got = G.gemm_tn_staged(ab, [b.data_ptr() for b, _ in bb], n_total, rows, ld, stage, snorm, flags, epoch, rank,
                                   epilogue=G.EPI_SQDIST, norm_a=na, peer_norm_ptrs=[nb.data_ptr() for _, nb in bb])
            torch.cuda.synchronize()
            assert (got - d2).abs().max().item() < 2e-3 * max(1.0, d2.abs().max().item())
    finally:
        G.set_variant(0)


@pytest.mark.skipif(__import__("os").environ.get("CML_EXPERIMENTAL") != "1", reason="opt-in until measured on hardware: set CML_EXPERIMENTAL=1")
@pytest.mark.parametrize("variant", [0, 1])
def test_staged_allgather_gemm_grouped_tile_order(variant):
    """Same check as above with the grouped tile order inside each source block switched on."""
    G.set_staged_grouping(True)
    try:
        test_staged_allgather_gemm_with_local_sources(8, 750, 2500, 3, variant)
        test_staged_allgather_gemm_with_local_sources(4, 1024, 512, 0, variant)
    finally:
        G.set_staged_grouping(False)
