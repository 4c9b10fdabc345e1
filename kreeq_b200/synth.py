"""Deterministic synthetic reads / assemblies of the BASELINE.json shapes, generated with
torch on the GPU (or CPU) and handed back as ASCII uint8 tensors in the layout the reference's
getKmers() builds: every read followed by 'N' (gfalibs/include/functions.h:112)."""
import torch

_ACGT = (65, 67, 71, 84)


def _lut(device):
    return torch.tensor(_ACGT, dtype=torch.uint8, device=device)


def random_genome_codes(n, seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return torch.randint(0, 4, (n,), dtype=torch.uint8, device=device, generator=g)


def codes_to_ascii(codes):
    return _lut(codes.device)[codes.long()]


def paired_reads_ascii(genome_codes, n_pairs, read_len, err, seed, insert_mean=400, insert_sd=50, chunk=1 << 20):
    """'N'-joined batch of 2*n_pairs reads (R1 forward strand, R2 reverse complement of the far end of
    the insert), substitution errors uniform over the 3 other bases.  Returns uint8 [2*n_pairs*(L+1)]."""
    dev = genome_codes.device
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    L = read_len
    G = genome_codes.numel()
    out = torch.empty((2 * n_pairs, L + 1), dtype=torch.uint8, device=dev)
    lut = _lut(dev)
    ar = torch.arange(L, device=dev)
    for p0 in range(0, n_pairs, chunk):
        n = min(chunk, n_pairs - p0)
        ins = (torch.randn(n, device=dev, generator=g) * insert_sd + insert_mean).clamp_(min=L).long()
        start = (torch.rand(n, device=dev, generator=g) * (G - ins.max().item() - 1)).long()
        flip = torch.rand(n, device=dev, generator=g) < 0.5          # which strand the fragment comes from
        r1 = genome_codes[(start[:, None] + ar[None, :])]
        r2 = 3 - genome_codes[(start + ins - 1)[:, None] - ar[None, :]]
        a = torch.where(flip[:, None], r2, r1)
        b = torch.where(flip[:, None], r1, r2)
        both = torch.stack((a, b), dim=1).reshape(2 * n, L)
        e = torch.rand(both.shape, device=dev, generator=g) < err
        shift = torch.randint(1, 4, both.shape, dtype=torch.uint8, device=dev, generator=g)
        both = torch.where(e, (both + shift) % 4, both)
        out[2 * p0:2 * (p0 + n), :L] = lut[both.long()]
    out[:, L] = 78                                                    # 'N'
    return out.reshape(-1)


def long_reads_ascii(genome_codes, n_reads, mean_len, sd_len, min_len, max_len, err, seed):
    """HiFi-like reads of varying length, 'N'-joined."""
    dev = genome_codes.device
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    G = genome_codes.numel()
    lens = (torch.randn(n_reads, device=dev, generator=g) * sd_len + mean_len).clamp_(min_len, max_len).long()
    lens = torch.minimum(lens, torch.tensor(G - 1, device=dev))
    start = (torch.rand(n_reads, device=dev, generator=g) * (G - lens).double()).long()
    offs = torch.cumsum(lens + 1, 0) - (lens + 1)
    total = int((lens + 1).sum().item())
    idx = torch.arange(total, device=dev)
    rid = torch.searchsorted(offs, idx, right=True) - 1
    within = idx - offs[rid]
    is_sep = within == lens[rid]
    flip = (torch.rand(n_reads, device=dev, generator=g) < 0.5)[rid]
    pos_f = start[rid] + within
    pos_r = start[rid] + lens[rid] - 1 - within
    pos = torch.where(flip, pos_r, pos_f).clamp_(0, G - 1)
    c = genome_codes[pos]
    c = torch.where(flip, 3 - c, c)
    e = torch.rand(total, device=dev, generator=g) < err
    shift = torch.randint(1, 4, (total,), dtype=torch.uint8, device=dev, generator=g)
    c = torch.where(e, (c + shift) % 4, c)
    out = _lut(dev)[c.long()]
    out[is_sep] = 78
    return out
