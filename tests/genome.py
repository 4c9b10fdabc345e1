"""Test-side restatement of how the reference turns a `-f` input into segments and paths, and of its
.kwig / .bed / .bkwig writers, used to check DBGbase arrays against reference outputs.

  loadGenome            src/input.cpp:188-308          (FASTA '>', FASTQ '@', else GFA)
  traverseInSequence    gfalibs/src/gfa.cpp:90-304     (split records at N/n/X/x runs into segments + gaps)
  printTableCompressed  src/kreeq-output.cpp:243-303   (.kwig)
  printTable            src/kreeq-output.cpp:138-241   (.bed)
  printTableCompressedBinary + writeIndex  src/kreeq-output.cpp:305-399 (.bkwig)
"""
import gzip
import struct

GAP_CHARS = b"NnXx"


def read_maybe_gz(path):
    with open(path, "rb") as f:
        data = f.read()
    return gzip.decompress(data) if data[:2] == b"\x1f\x8b" else data


def split_record(seq):
    """-> components [("S", start, bytes) | ("G", start, length)] of one record (newlines removed)."""
    comps, i, n = [], 0, len(seq)
    while i < n:
        j = i
        if seq[i] in GAP_CHARS:
            while j < n and seq[j] in GAP_CHARS:
                j += 1
            comps.append(("G", i, j - i))
        else:
            while j < n and seq[j] not in GAP_CHARS:
                j += 1
            comps.append(("S", i, seq[i:j]))
        i = j
    return comps


def load_paths(data):
    """-> [(header, components)] in file order."""
    paths = []
    if data[:1] == b">":
        pos = 1                             # header = rest of the line ('>' allowed inside), sequence = up to the next '>'
        while pos < len(data):
            nl = data.find(b"\n", pos)
            nl = len(data) if nl < 0 else nl
            head = data[pos:nl]
            nxt = data.find(b">", nl)
            nxt = len(data) if nxt < 0 else nxt
            body = data[nl + 1:nxt]
            paths.append((head.split(b" ")[0].decode(), split_record(body.replace(b"\n", b""))))
            pos = nxt + 1
    elif data[:1] == b"@":
        lines = data.split(b"\n")
        for i in range(0, len(lines) - 1, 4):
            if not lines[i]:
                break
            paths.append((lines[i][1:].split(b" ")[0].decode(), split_record(lines[i + 1])))
    else:                                   # GFA: S lines are segments, P lines name single-segment paths
        segs = {}
        for ln in data.split(b"\n"):
            f = ln.split(b"\t")
            if f[0] == b"S":
                segs[f[1]] = f[2]
        named = False
        for ln in data.split(b"\n"):
            f = ln.split(b"\t")
            if f[0] == b"P":
                named = True
                comps, pos = [], 0
                for s in f[2].split(b","):
                    seq = segs[s[:-1]]
                    comps.append(("S", pos, seq)); pos += len(seq)
                paths.append((f[1].decode(), comps))
        if not named:
            for name, seq in segs.items():
                paths.append((name.decode(), [("S", 0, seq)]))
    return paths


def segments(paths):
    return [c[2] for _, comps in paths for c in comps if c[0] == "S"]


def _fb(b):
    """(cov, F, B) of one DBGbase entry as the writers print it (kreeq-output.cpp:277, 379-383)."""
    fw, bw, cov, is_fw = int(b["fw"]), int(b["bw"]), int(b["cov"]), bool(b["isFw"])
    return cov, (fw if is_fw else bw), (bw if is_fw else fw)


def kwig_text(k, paths, bases):
    """bases: list of DBGbase arrays, one per segment in segments(paths) order."""
    out = ["%d\n" % k]
    it = iter(bases)
    for head, comps in paths:
        pos = 0
        for c in comps:
            if c[0] == "S":
                arr = next(it)
                out.append("fixedStep chrom=%s start=%d step=1\n" % (head, pos))
                for b in arr:
                    out.append("%d,%d,%d\n" % _fb(b))
                pos += len(arr)
            else:
                pos += c[2]
    return "".join(out).encode()


def bed_text(k, paths, bases):
    """printTable: sliding window of the last k (cov, F, B) triples, zeros before the segment start."""
    out = []
    it = iter(bases)
    for head, comps in paths:
        pos = 0
        for c in comps:
            if c[0] == "S":
                arr = next(it)
                cv, fv, bv = [0] * k, [0] * k, [0] * k
                for b in arr:
                    cov, F, B = _fb(b)
                    cv = cv[1:] + [cov]; fv = fv[1:] + [F]; bv = bv[1:] + [B]
                    out.append("%s\t%d\t%s\t%s\t%s\n" % (head, pos, ":".join(map(str, cv)), ":".join(map(str, fv)),
                                                        ":".join(map(str, bv))))
                    pos += 1
            else:
                pos += c[2]
    return "".join(out).encode()


def bkwig_bytes(k, paths, bases):
    out = [struct.pack("<B", k)]
    idx = [struct.pack("<I", len(paths))]
    for head, comps in paths:
        segs = [c for c in comps if c[0] == "S"]
        idx.append(struct.pack("<H", len(head)) + head.encode() + struct.pack("<I", len(segs)))
        for c in segs:
            idx.append(struct.pack("<QQB", c[1], len(c[2]), 1))
    out += idx
    for arr in bases:
        for b in arr:
            out.append(struct.pack("<III", *_fb(b)))
    return b"".join(out)
