"""CPU oracle for the two hot paths.  TEST INFRASTRUCTURE ONLY.

Nothing in the product package (`simclr_pytorch_distributed_b200/`) may import this
package.  The only legal users are `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py`, and there only as the checker
or as the timed CPU baseline -- never as the thing shipped.

Parity status: the reference (`/root/reference`) ships no tests, golden vectors or
fixtures, so the oracle is pinned against *outputs of the reference
itself*: `tests/golden/make_golden.py` imports `/root/reference/losses.py` unmodified (and
`torch.nn.BatchNorm2d` on the rank-concatenated batch for SyncBN, whose CUDA-only ATen ops
cannot run on a CPU) and stores input/output vectors under `tests/golden/`;
`tests/test_oracle_golden.py` checks every function here against them.
"""
from .supcon_oracle import (  # noqa: F401
    supcon_forward_backward,
    supcon_row_stats,
    gather_glue,
    normalize_rows,
    normalize_rows_backward,
)
from .syncbn_oracle import (  # noqa: F401
    bn_local_stats,
    bn_merge_stats,
    syncbn_forward,
    syncbn_backward,
)
