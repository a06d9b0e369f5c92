"""Runs pytest in-process after filling the caching allocator's free blocks with NaN: a kernel that reads memory nobody wrote
(an output allocated with torch.empty and only partly written, a scratch buffer) then fails deterministically instead of once in
a while.  usage: python tools/poison_run.py [GB] <pytest args>"""
import sys

import pytest
import torch

gb = float(sys.argv[1])
for _ in range(2):
    blocks = [torch.full((int(gb * 1e9 / 4 / 8),), float('nan'), device='cuda') for _ in range(8)]
    del blocks
torch.cuda.synchronize()
sys.exit(pytest.main(sys.argv[2:]))
