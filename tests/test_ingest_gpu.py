"""GPU parity of the ingest kernels (normalise+pack, masked mean-pool+normalise+pack) through the C ABI."""
import numpy as np
import pytest

from oracle import ref_ingest
from helpers import load

pytestmark = pytest.mark.gpu


def _unpack(rows, dtype):
    return ref_ingest.unpack_bf16(rows) if dtype == "bf16" else rows.astype(np.float32)


def _pack(x, dtype):
    if dtype == "fp16":
        return ref_ingest.pack_f16(x).astype(np.float32)
    if dtype == "bf16":
        return ref_ingest.unpack_bf16(ref_ingest.pack_bf16(x))
    return x.astype(np.float32)


ULP = {"fp16": 2.0 ** -11, "bf16": 2.0 ** -8, "fp32": 2.0 ** -23}


def _assert_packed(got, want_f32, dtype):
    """Equal to the oracle's packed value up to one unit in the last place of the packed type
    (the fp32 norm may differ in its last bit with summation order), and mostly bit-equal.
    For an fp32 store every quotient inherits the norm's last-bit difference, so only the
    2-ulp bound is asserted there."""
    want = _pack(want_f32, dtype)
    diff = np.abs(got - want)
    assert np.all(diff <= 2 * ULP[dtype] * np.maximum(np.abs(want), 2.0 ** -14)), float(diff.max())
    if dtype != "fp32":
        assert np.mean(got == want) > 0.98


@pytest.mark.parametrize("dtype", ["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("d", [128, 384])
def test_norm_pack_golden(d, dtype):
    """Fixtures made by executing ollama_rag_system.py:379-380 on numpy."""
    from modular_rag_system_b200 import DeviceStore
    fx = load(f"normpack_d{d}.npz")
    s = DeviceStore(d, dtype, 64)
    assert s.add(fx["x"]) == 0
    rows, norms = s.get_rows()
    np.testing.assert_allclose(norms, fx["norms"], rtol=3e-7)
    _assert_packed(_unpack(rows, dtype), fx["unit"], dtype)


@pytest.mark.parametrize("dim", [1, 7, 100, 384, 1000])
def test_norm_pack_shapes_and_zero_rows(dim):
    from modular_rag_system_b200 import DeviceStore
    rng = np.random.default_rng(dim)
    x = (rng.standard_normal((37, dim)) * rng.uniform(1e-3, 50, (37, 1))).astype(np.float32)
    x[3] = 0
    s = DeviceStore(dim, "fp16", 8)  # forces a reserve
    s.add(x[:5]); s.add(x[5:])
    rows, norms = s.get_rows()
    unit, onorm = ref_ingest.normalize_f32(x)
    np.testing.assert_allclose(norms, onorm, rtol=3e-7)
    _assert_packed(rows.astype(np.float32), unit, "fp16")
    assert (rows[3] == 0).all() and norms[3] == 0


def test_add_dev_from_torch():
    import torch
    from modular_rag_system_b200 import DeviceStore
    rng = np.random.default_rng(0)
    x = rng.standard_normal((300, 384)).astype(np.float32)
    for tdt in (torch.float32, torch.float16, torch.bfloat16):
        s = DeviceStore(384, "fp16", 300)
        xt = torch.from_numpy(x).cuda().to(tdt)
        s.add_dev(xt)
        torch.cuda.synchronize()
        rows, norms = s.get_rows()
        unit, onorm = ref_ingest.normalize_f32(xt.float().cpu().numpy())
        np.testing.assert_allclose(norms, onorm, rtol=3e-7)
        _assert_packed(rows.astype(np.float32), unit, "fp16")


@pytest.mark.parametrize("tok_dtype", ["fp32", "fp16", "bf16"])
@pytest.mark.parametrize("B,S,H", [(9, 256, 384), (5, 37, 100), (3, 512, 768), (2, 16, 2304), (4, 1, 8)])
def test_pool_norm_pack_vs_oracle(B, S, H, tok_dtype):
    """PARITY UNPINNED (sentence-transformers is un-vendored): checked against the restated
    Pooling(mean)+Normalize of oracle/ref_ingest.py."""
    from modular_rag_system_b200 import DeviceStore
    rng = np.random.default_rng(B * S + H)
    tok = rng.standard_normal((B, S, H)).astype(np.float32)
    lens = rng.integers(1, S + 1, size=B)
    mask = (np.arange(S)[None, :] < lens[:, None])
    if tok_dtype == "fp16":
        tok_in = tok.astype(np.float16); tok_ref = tok_in.astype(np.float32)
    elif tok_dtype == "bf16":
        tok_in = ref_ingest.pack_bf16(tok); tok_ref = ref_ingest.unpack_bf16(tok_in)
    else:
        tok_in = tok_ref = tok
    unit, onorm = ref_ingest.pool_normalize_f32(tok_ref, mask)
    for mdt in (np.uint8, np.int64, np.float32, np.int32):
        s = DeviceStore(H, "fp16", B)
        s.add_pooled(tok_in, mask.astype(mdt))
        rows, norms = s.get_rows()
        np.testing.assert_allclose(norms, onorm, rtol=2e-5)
        got = rows.astype(np.float32)
        assert np.abs(got - unit).max() <= 2.0 ** -11 + 1e-5
        assert np.mean(got == ref_ingest.pack_f16(unit).astype(np.float32)) > 0.97


def test_pool_edge_cases():
    from modular_rag_system_b200 import DeviceStore
    rng = np.random.default_rng(1)
    tok = rng.standard_normal((4, 12, 64)).astype(np.float32)
    mask = np.ones((4, 12), np.int64)
    mask[1] = 0              # fully masked chunk -> zero vector (sum 0 / 1e-9, then / 1e-12)
    mask[2, 1:] = 0          # single valid token
    tok[3, 5:] = np.nan; mask[3, 5:] = 0   # masked tokens are never read
    s = DeviceStore(64, "fp32", 4)
    s.add_pooled(tok, mask)
    rows, norms = s.get_rows()
    unit, onorm = ref_ingest.pool_normalize_f32(np.nan_to_num(tok), mask)
    np.testing.assert_allclose(rows, unit, atol=2e-6)
    assert (rows[1] == 0).all() and norms[1] == 0
    np.testing.assert_allclose(rows[2], tok[2, 0] / np.linalg.norm(tok[2, 0]), atol=2e-6)
    # no mask = plain mean
    s2 = DeviceStore(64, "fp32", 4)
    s2.add_pooled(np.nan_to_num(tok), None)
    r2, _ = s2.get_rows()
    u2, _ = ref_ingest.pool_normalize_f32(np.nan_to_num(tok), np.ones((4, 12)))
    np.testing.assert_allclose(r2, u2, atol=2e-6)


def test_pooled_rows_are_searchable():
    import torch
    from modular_rag_system_b200 import DeviceStore
    g = torch.Generator(device="cuda").manual_seed(2)
    tok = torch.randn((64, 32, 384), generator=g, device="cuda", dtype=torch.float16)
    mask = torch.ones((64, 32), dtype=torch.uint8, device="cuda")
    s = DeviceStore(384, "fp16", 64)
    s.add_pooled_dev(tok, mask)
    torch.cuda.synchronize()
    q = tok[17].float().mean(0).cpu().numpy()
    ids, sc = s.search(q, 1, "cosine")
    assert ids[0][0] == 17 and sc[0][0] > 0.999


def test_pool_full_size_slab_properties():
    """BASELINE configs[2] shape (seq 256, hidden 384, fp16 tokens, ragged u8 masks), one 16 384-chunk slab of the
    1M: every packed row is a unit vector, the stored norm is the norm of the masked mean, and the result equals a
    torch fp32 statement of Pooling(mean) + Normalize on the same device tensors."""
    import torch
    from modular_rag_system_b200 import DeviceStore
    B, S, H = 16384, 256, 384
    g = torch.Generator(device="cuda").manual_seed(2)
    tok = torch.randn((B, S, H), generator=g, device="cuda", dtype=torch.float16)
    lens = torch.randint(16, S + 1, (B, 1), generator=g, device="cuda")
    mask = (torch.arange(S, device="cuda")[None, :] < lens).to(torch.uint8).contiguous()
    st = DeviceStore(H, "fp16", B)
    st.add_pooled_dev(tok, mask)
    torch.cuda.synchronize()
    rows, norms = st.get_rows(0, B)
    m = mask.float()
    mean = (tok.float() * m[:, :, None]).sum(1) / m.sum(1, keepdim=True).clamp(min=1e-9)
    ref = torch.nn.functional.normalize(mean, p=2, dim=1, eps=1e-12)
    got = torch.from_numpy(rows.astype(np.float32)).cuda()
    assert (got - ref).abs().max().item() < 1e-3                      # one fp16 rounding of a unit vector's entry
    assert (got.norm(dim=1) - 1).abs().max().item() < 2e-3
    assert np.allclose(norms, mean.norm(dim=1).cpu().numpy(), rtol=1e-4)
    st.close()
