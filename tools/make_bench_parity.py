#!/usr/bin/env python
"""Writes tests/golden/bench_parity.json: the answers of the seeded parity problems of kreeq_b200/paritycheck.py,
computed by the CPU oracle (oracle/kreeq_oracle.c, pinned to the reference by tests/test_oracle_golden.py).
bench.py and __graft_entry__.smoke() compare the GPU path with these constants at every GPU count.

    python tools/make_bench_parity.py
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import numpy as np          # noqa: E402
import kreeq_oracle as ko   # noqa: E402
from kreeq_b200 import paritycheck as pc   # noqa: E402

out = {}
for name in pc.PROBLEMS:
    t0 = time.time()
    reads, genome, k = pc.problem(name)
    tab = ko.Table(k)
    inst = tab.count_batch(reads.tobytes())
    keys, recs = tab.export()
    st = tab.stats()
    n_contigs = 4
    clen = genome.size // n_contigs
    acc = [0, 0, 0]
    rd = 0
    for i in range(n_contigs):
        seg = genome[i * clen:(i + 1) * clen].tobytes()
        rc, rows, a = tab.validate_segment(seg)
        assert rc == 0
        acc = [x + y for x, y in zip(acc, a)]
        rd = pc.add_mod64(rd, pc.rows_digest(rows, i * clen))
    out[name] = {"instances": int(inst), "db": pc.db_digest(keys, recs), "stats": {k_: int(v) for k_, v in st.items()} if isinstance(st, dict) else None,
                 "contigs": n_contigs, "qv": [int(x) for x in acc], "rows": rd}
    print(name, out[name], "%.1f s" % (time.time() - t0))
with open(pc.GOLDEN, "w") as f:
    json.dump(out, f, indent=1)
