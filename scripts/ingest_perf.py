"""Ingest kernels timing (development aid): pool+normalise+pack and normalise+pack."""
import sys, torch
sys.path.insert(0, ".")
import modular_rag_system_b200 as pkg
S, H, slab = 256, 384, 16384
g = torch.Generator(device="cuda").manual_seed(2)
for dt in (torch.float16, torch.float32):
    tok = torch.randn((slab if dt == torch.float16 else slab // 2, S, H), generator=g, device="cuda", dtype=dt)
    B = tok.shape[0]
    lens = torch.randint(16, S + 1, (B, 1), generator=g, device="cuda")
    for name, mask in (("ragged", (torch.arange(S, device="cuda")[None, :] < lens).to(torch.uint8).contiguous()),
                       ("full", torch.ones((B, S), dtype=torch.uint8, device="cuda"))):
        st = pkg.DeviceStore(H, "fp16", 40 * B)
        for _ in range(3): st.add_pooled_dev(tok, mask)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(20): st.add_pooled_dev(tok, mask)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        es = tok.element_size()
        full = B * (S * H * es + S + H * 2 + 4)
        valid = float(mask.sum().item()) * H * es + B * (S + H * 2 + 4)
        print(f"pool {dt} {name:7s}: {ms:7.3f} ms  {B/ms*1e3/1e6:6.2f} M chunks/s  nominal {full/ms/1e6:7.0f} GB/s  actually-read {valid/ms/1e6:7.0f} GB/s", flush=True)
        st.close()
X = torch.randn((2_000_000, 384), generator=g, device="cuda")
st = pkg.DeviceStore(384, "fp16", 12_000_000)
st.add_dev(X); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(5): st.add_dev(X)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"norm_pack fp32->fp16 2M x 384: {ms:.3f} ms  {2e6*(384*4+384*2+4)/ms/1e6:.0f} GB/s")
