"""norm_pack (K2) timing (development aid): python scripts/norm_pack_perf.py
Set RAGB_NORM_PACK_GENERIC=1 for the two-pass generic kernel."""
import os, sys, torch
sys.path.insert(0, ".")
import modular_rag_system_b200 as pkg
g = torch.Generator(device="cuda").manual_seed(2)
tag = "generic" if os.environ.get("RAGB_NORM_PACK_GENERIC") else "reg"
for n, d, din, dout in ((2_000_000, 384, torch.float32, "fp16"), (1_000_000, 768, torch.float32, "fp16"),
                        (2_000_000, 384, torch.float16, "fp16"), (1_000_000, 1024, torch.float32, "bf16"),
                        (1_000_000, 384, torch.float32, "fp32")):
    X = torch.randn((n, d), generator=g, device="cuda").to(din)
    st = pkg.DeviceStore(d, dout, 12 * n)
    st.add_dev(X); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10): st.add_dev(X)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    ob = {"fp16": 2, "bf16": 2, "fp32": 4}[dout]
    print(f"norm_pack[{tag}] {str(din)[6:]}->{dout} {n} x {d}: {ms:.3f} ms  {n * (d * X.element_size() + d * ob + 4) / ms / 1e6:.0f} GB/s", flush=True)
    st.close(); del X, st
    torch.cuda.empty_cache()
