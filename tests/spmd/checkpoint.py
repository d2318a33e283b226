# checkpoint / resume through rank-0 save + synchronise-on-load; trace regions.
import sys
import torch
import fluxmpi_b200 as fm
from fluxmpi_b200.utils import checkpoint, trace

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
path = sys.argv[1]

tree = {"w": torch.full((3, 2), float(r + 1)), "step": 7 + r, "nested": (torch.arange(4.0) * (r + 1), "tag")}
checkpoint.save(path, tree, root=0)                       # rank 0's values land on disk
mine = {"w": torch.zeros(3, 2), "step": 0, "nested": (torch.zeros(4), "tag")}
with trace.region("load"):
    got = checkpoint.load(path, mine, root=0)
assert torch.equal(got["w"], torch.ones(3, 2)) and got["w"] is mine["w"]
assert got["step"] == 7 and torch.equal(got["nested"][0], torch.arange(4.0)) and got["nested"][1] == "tag"
s = trace.summary()
assert s["load"]["count"] == 1 and s["load"]["total_ms"] >= 0
trace.report()
# collective counters: calls and bytes per public collective, host and device side
trace.reset()
fm.allreduce_(torch.ones(10), "sum")
fm.allreduce_(torch.ones(6, dtype=torch.float64), "max")
fm.bcast_(torch.ones(4), 0)
fm.allreduce_gradients({"a": torch.ones(3), "b": torch.ones(5)})
c = trace.counters()
assert c["allreduce/host"] == {"calls": 2, "bytes": 40 + 48}, c
assert c["bcast/host"] == {"calls": 1, "bytes": 16}
assert c["allreduce_tree/host"] == ({"calls": 1, "bytes": 32} if w > 1 else c["allreduce_tree/host"])
fm.Finalize()
print("RANK_OK", r)
