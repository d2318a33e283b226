# mirrors /root/reference/test/test_data.jl:1-29
import math
import torch
import fluxmpi_b200 as fm

fm.Init(verbose=True)
r, w = fm.local_rank(), fm.total_workers()

data = torch.randn(10, generator=torch.Generator().manual_seed(0))
tot = float(data.double().sum())
ddc = fm.DistributedDataContainer(list(data))
per = math.ceil(10 / w)
if r != w - 1:
    assert len(ddc) == per
else:
    assert len(ddc) == 10 - (w - 1) * per
# contiguous blocks in rank order (src/data.jl:18-19)
assert list(ddc.idxs) == list(range(r * per, min(r * per + per, 10)))

dsum = torch.zeros(1, dtype=torch.float64)
for i in range(len(ddc)):
    dsum += ddc[i].double()
fm.allreduce_(dsum, "+")
assert abs(float(dsum) - tot) < 1e-9

# N=9, W=4: the reference throws; here trailing ranks get an empty shard
sizes = [len(fm.DistributedDataContainer(list(range(9)), rank=q, world=4)) for q in range(4)]
assert sizes == [3, 3, 3, 0]

# loader over the shard (CPU path)
xs = torch.arange(40, dtype=torch.float32).reshape(20, 2)
ys = torch.arange(20)
shard = fm.DistributedDataContainer(fm.TensorDataset(xs, ys))
seen = []
for xb, yb in fm.DeviceLoader(shard, batch_size=2, device="cpu", drop_last=False):
    assert xb.shape[1] == 2 and xb.shape[0] == yb.shape[0]
    seen += yb.tolist()
assert seen == list(shard.idxs)
cnt = torch.tensor([len(seen)])
fm.allreduce_(cnt, "sum")
assert int(cnt) == 20
fm.Finalize()
print("RANK_OK", r)
