"""Count one C3 / 16 read set (6.25 Mbp genome, 26x) with and without planted repeats (64 x 5 kb) and a 2 kb poly-A:
wall time and the kernel timers (0 pipeline, 2 insert, 3 level 1, 4 level 2) of the second pass over the same reads.
Shows what runs of one key cost the slice-ordered insert."""
import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import kreeq_b200 as kb
from kreeq_b200 import synth
dev = torch.device("cuda", 0)
k, L, err = 21, 150, 0.01
gsize, pairs = 6_250_000, 543_750
for name, nrep, polya in (("plain", 0, 0), ("repeats", 64, 0), ("polyA", 0, 2000), ("both", 64, 2000)):
    genome = synth.random_genome_codes(gsize, 2, dev)
    rng = np.random.default_rng(3)
    rep = torch.from_numpy(rng.integers(0, 4, 5000).astype(np.uint8)).to(dev)
    places = rng.choice(gsize // 7000 - 1, 65, replace=False) * 7000
    for s in places[:nrep]:
        genome[int(s):int(s) + 5000] = rep
    if polya: genome[int(places[64]):int(places[64]) + polya] = 0
    reads = synth.paired_reads_ascii(genome, pairs, L, err, 30).cpu().numpy()
    est = int((gsize + 2 * pairs * L * err * k * 0.85) * 1.05)
    with kb.DBG(k=k, expected_distinct=est) as g:
        g.traverseInReads(reads); g.finalize(); g.clear()
        g.set_timing(True)
        for w in range(8): g.kernel_time_ms(w)
        t = time.perf_counter(); g.traverseInReads(reads); g.finalize(); g.sync(); dt = time.perf_counter() - t
        print(name, "wall %.1f ms" % (dt * 1e3), {w: round(g.kernel_time_ms(w)[0], 2) for w in (0, 2, 3, 4)}, flush=True)
