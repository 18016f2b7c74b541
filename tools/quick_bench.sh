#!/bin/bash
# tuning runs: the resident leg and per-kernel times of the given configurations (default C4 C2 C3 C5), one JSON line each
for c in ${@:-C4 C2 C3 C5}; do python bench.py --resident-only --config $c --steps 5 --warmup 3 2>&1 | tail -1; done
