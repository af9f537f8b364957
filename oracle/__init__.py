"""TEST INFRASTRUCTURE ONLY.

CPU restatement of baby-sglang's paged-attention hot path.  Nothing in the
product package (`baby_sglang_b200/`) imports this directory.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may use it, and only as the checker or the timed CPU baseline.
"""
