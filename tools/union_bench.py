"""Prints the database-row timings (save / union / load) of one GPU as one JSON line; see kreeq_b200/dbbench.py."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-mbp", type=float, default=10.0)
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--dir", default=None)
    a = ap.parse_args()
    from kreeq_b200.dbbench import union_rows
    print(json.dumps(union_rows(a.genome_mbp, a.coverage, a.dir)))
