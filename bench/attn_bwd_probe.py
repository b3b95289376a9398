"""Timing probe for the attention kernels: `python bench/attn_bwd_probe.py TAG [B H T]` prints the median device time
of one backward (memset + delta prep + main + finish) and one forward, L2 flushed before each call.  Used with
DTB_ATTN_BWD_WS / DTB_ATTN_DBG (timing experiments that disable parts of the warp-specialised backward)."""
import statistics
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from distributed_training_b200.ops import native  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else ""
B, H, T = (int(v) for v in sys.argv[2:5]) if len(sys.argv) >= 5 else (8, 12, 1024)
C = H * 64
qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
dout = torch.randn(B, T, C, device="cuda").to(torch.bfloat16)
out = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device="cuda")
dqkv = torch.empty_like(qkv)
scr = torch.empty(native.attention_bwd_scratch_floats(B, T, H), device="cuda")
native.attention_fwd(qkv, out, lse, H)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def med(fn):
    ts = []
    for _ in range(15):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    return round(statistics.median(ts[5:]), 1)


print(tag, f"B{B} H{H} T{T}", "bwd_us", med(lambda: native.attention_bwd(qkv, out, dout, lse, dqkv, scr, H)),
      "fwd_us", med(lambda: native.attention_fwd(qkv, out, lse, H)))
