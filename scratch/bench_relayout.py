"""Time of the ResNet-18 filter-copy refresh (one iit_filter_relayout_batched launch) (development tool)."""
import sys, torch
sys.path.insert(0, ".")
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device="cuda", dtype=torch.bfloat16)
ll.refresh_filter_shadow(); torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for _ in range(7):
    flush.zero_()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); ll.refresh_filter_shadow(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
print("refresh_filter_shadow", sorted(ts)[3] * 1e3, "us")
