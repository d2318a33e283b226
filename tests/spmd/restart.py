# A training loop that checkpoints every step and crashes once: with `launch --max-restarts 1`
# the job is torn down, started again, and continues from the last save instead of from zero.
import os
import sys

import torch

import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R
from fluxmpi_b200.utils import checkpoint

path = sys.argv[1]
attempt = int(os.environ["FMPI_RESTART_COUNT"])
fm.Init()
r, w = fm.local_rank(), fm.total_workers()

ps = {"w": torch.zeros(4)}
st = R.setup(fm.DistributedOptimizer(R.Descent(1.0)), ps)
state = {"ps": ps, "step": 0}
if os.path.exists(path):                                   # resume
    state = checkpoint.load(path, state)
start = int(state["step"])
assert (attempt == 0) == (start == 0), (attempt, start)

for step in range(start, 6):
    gs = {"w": torch.full((4,), -1.0)}                     # w grows by W per step (gradients are summed)
    st, ps = R.update_(st, state["ps"], gs)
    state = {"ps": ps, "step": step + 1}
    checkpoint.save(path, state)
    if attempt == 0 and step == 2 and r == w - 1:
        sys.stdout.flush()
        os._exit(7)                                        # a rank dies after three saved steps

assert torch.equal(state["ps"]["w"], torch.full((4,), 6.0 * w)), state["ps"]["w"]
assert start == 3, start                                   # the second attempt did steps 3..5 only
fm.Finalize()
print(f"RANK_OK {r} resumed_from={start}", flush=True)
