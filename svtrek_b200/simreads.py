"""Seeded synthetic long-read sets with planted DEL/INS, in the SoA layout the B200 path consumes.

This is the input generator for BASELINE.json's five configs: reads are produced
directly as packed BAM-native CIGAR words (len<<4 | op) sorted by (tid, pos) -- the same order a
coordinate-sorted BAM yields -- so the full-size configs never have to pass through BGZF.  Small sets can
also be written as SAM (+ random SEQ/QUAL so the BAM compresses realistically) for the reference CLI.

VCF lines are written in the exact shape the reference's utils/simvcf.py produces
(``...;CIPOS=a,b;CIEND=c,d;END=e;SVELDT=SIMULATED;SVTYPE=DEL`` -- simvcf.py:173-186), plus the
"END-first" twin that makes the reference's strstr("END=") find the real END (audit.c:145).
Nothing here is on the product path; bench.py and the tests use it to make inputs.
"""
from __future__ import annotations

import dataclasses
import numpy as np

OP_M, OP_I, OP_D, OP_N, OP_S, OP_H, OP_P, OP_EQ, OP_X = range(9)
OPCHARS = "MIDNSHP=X"
JITTER = np.array([0, 0, 0, 0, 0, 1, -1, 2, -2, 3], dtype=np.int64)


@dataclasses.dataclass
class ReadSet:
    """SoA long-read set: reads sorted by (tid, pos). cigar words are BAM-native: len<<4 | op."""
    tid_off: np.ndarray      # int64[n_tid+1]  read index range per contig
    pos: np.ndarray          # uint32[n_reads] 0-based leftmost position (bam1_core_t.pos)
    end: np.ndarray          # uint32[n_reads] bam_endpos(): pos + max(1, ref-consumed length)
    n_cigar: np.ndarray      # uint32[n_reads]
    cigar_off: np.ndarray    # uint64[n_reads+1] word offset of each read's CIGAR
    cigar: np.ndarray        # uint32[n_words]
    contig_len: list

    @property
    def n_reads(self) -> int:
        return int(self.pos.shape[0])

    @property
    def n_tid(self) -> int:
        return int(self.tid_off.shape[0] - 1)


@dataclasses.dataclass
class SVTruth:
    tid: np.ndarray    # int32
    pos0: np.ndarray   # int64 0-based breakpoint (first deleted base / insertion point)
    length: np.ndarray  # int64
    is_del: np.ndarray  # bool


def ref_consumed(words: np.ndarray) -> np.ndarray:
    op = words & 15
    ln = (words >> 4).astype(np.int64)
    # M, D, N, =, X consume the reference (bam_cigar_type & 2)
    consumes = (op == OP_M) | (op == OP_D) | (op == OP_N) | (op == OP_EQ) | (op == OP_X)
    return np.where(consumes, ln, 0)


def endpos_from_cigar(pos, cigar_off, cigar) -> np.ndarray:
    """bam_endpos (htslib sam.c:638-643) for every read, vectorised."""
    n = pos.shape[0]
    rc = ref_consumed(cigar)
    csum = np.zeros(cigar.shape[0] + 1, dtype=np.int64)
    np.cumsum(rc, out=csum[1:])
    rlen = csum[cigar_off[1:].astype(np.int64)] - csum[cigar_off[:-1].astype(np.int64)]
    rlen = np.where(rlen == 0, 1, rlen)
    return (pos.astype(np.int64) + rlen).astype(np.uint32)[:n]


def plant_svs(rng, contig_len, n_sv, *, min_len=60, max_len=2000, min_gap=1500, margin=25000,
              kind="alt", log_uniform=False) -> tuple:
    """SV breakpoints on one contig: sorted, >= min_gap apart (and clear of the previous DEL's span)."""
    span = contig_len - 2 * margin
    if n_sv <= 0 or span <= 0:
        z = np.zeros(0, dtype=np.int64)
        return z, z, np.zeros(0, dtype=bool)
    if log_uniform:
        length = np.exp(rng.uniform(np.log(min_len), np.log(max_len), n_sv)).astype(np.int64)
    else:
        length = rng.integers(min_len, max_len + 1, n_sv)
    if kind == "alt":
        is_del = (np.arange(n_sv) % 2) == 0
    elif kind == "ins":
        is_del = np.zeros(n_sv, dtype=bool)
    else:
        is_del = np.ones(n_sv, dtype=bool)
    need = np.maximum(np.where(is_del, length + 200, 0), min_gap)   # room each SV takes before the next may start
    slack = span - int(need.sum())
    if slack < 0:
        raise ValueError("contig too small for the requested SV density")
    cuts = np.sort(rng.integers(0, slack + 1, n_sv))
    pos0 = margin + cuts + np.concatenate(([0], np.cumsum(need)[:-1]))
    return pos0.astype(np.int64), length.astype(np.int64), is_del


def _expand_ranges(lo, hi):
    """flat (owner, index) pairs for integer ranges [lo_i, hi_i)."""
    cnt = np.maximum(hi - lo, 0)
    owner = np.repeat(np.arange(lo.shape[0]), cnt)
    start = np.repeat(lo, cnt)
    base = np.repeat(np.cumsum(cnt) - cnt, cnt)
    idx = start + (np.arange(int(cnt.sum())) - base)
    return owner, idx


def gen_contig(rng, contig_len, n_reads, read_len, small_rate, sv_pos0, sv_len, sv_is_del, *,
               carrier_frac=0.8, split_frac=0.05, clip_len=(200, 3000), chunk=1 << 17):
    """One contig's reads -> (pos, n_cigar, cigar words). Vectorised, chunked over reads."""
    starts_all = np.sort(rng.integers(0, max(1, contig_len - read_len), n_reads)).astype(np.int64)
    out_pos, out_n, out_words = [], [], []
    for c0 in range(0, n_reads, chunk):
        s = starts_all[c0:c0 + chunk].copy()
        nr = s.shape[0]
        e = s + read_len
        lead = np.zeros(nr, dtype=np.int64)   # leading soft clip length (0 = none)
        trail = np.zeros(nr, dtype=np.int64)

        # --- planted SVs overlapped by each read (breakpoint well inside the read)
        lo = np.searchsorted(sv_pos0, s + 50, side="left")
        hi = np.searchsorted(sv_pos0, e - 50, side="left")
        rid_sv, k = _expand_ranges(lo, hi)
        carried = rng.random(rid_sv.shape[0]) < carrier_frac
        rid_sv, k = rid_sv[carried], k[carried]
        bp = sv_pos0[k] + JITTER[rng.integers(0, JITTER.shape[0], k.shape[0])]
        ln = sv_len[k]
        isd = sv_is_del[k]
        # split reads: first chosen SV per read cuts the read and becomes a soft clip
        is_split = rng.random(rid_sv.shape[0]) < split_frac
        if is_split.any():
            cand = np.flatnonzero(is_split)
            first = np.unique(rid_sv[cand], return_index=True)[1]
            cand = cand[first]
            left_part = rng.random(cand.shape[0]) < 0.5
            cl = rng.integers(clip_len[0], clip_len[1] + 1, cand.shape[0])
            r = rid_sv[cand]
            # A: keep [s, bp) + trailing clip ; B: leading clip + [bp(+len), e)
            e[r[left_part]] = bp[cand[left_part]]
            trail[r[left_part]] = cl[left_part]
            nb = np.where(isd[cand[~left_part]], bp[cand[~left_part]] + ln[cand[~left_part]], bp[cand[~left_part]])
            rb, clb = r[~left_part], cl[~left_part]
            fits = nb <= e[rb] - 50                          # a deletion reaching past the read end cannot be a B split
            s[rb[fits]] = nb[fits]
            lead[rb[fits]] = clb[fits]
        ev_rid = [rid_sv]
        ev_pos = [bp]
        ev_op = [np.where(isd, OP_D, OP_I)]
        ev_len = [ln]

        # --- small indels (sequencing noise)
        if small_rate > 0:
            cnt = rng.poisson(small_rate * read_len, nr)
            rid_sm = np.repeat(np.arange(nr), cnt)
            p = starts_all[c0:c0 + chunk][rid_sm] + rng.integers(1, read_len - 1, rid_sm.shape[0])
            near = np.searchsorted(sv_pos0, p + 60, side="right") - 1       # last SV starting <= p+60
            near = np.clip(near, 0, max(0, sv_pos0.shape[0] - 1))
            if sv_pos0.shape[0]:
                blocked = (p >= sv_pos0[near] - 60) & (p <= sv_pos0[near] + np.where(sv_is_del[near], sv_len[near], 0) + 60)
            else:
                blocked = np.zeros(p.shape[0], dtype=bool)
            keep = ~blocked
            rid_sm, p = rid_sm[keep], p[keep]
            sm_del = rng.random(p.shape[0]) < 0.5
            sm_len = 1 + (rng.random(p.shape[0]) < 0.15) + (rng.random(p.shape[0]) < 0.03)
            ev_rid.append(rid_sm); ev_pos.append(p); ev_op.append(np.where(sm_del, OP_D, OP_I)); ev_len.append(sm_len.astype(np.int64))

        rid = np.concatenate(ev_rid); p = np.concatenate(ev_pos)
        op = np.concatenate(ev_op).astype(np.int64); ln = np.concatenate(ev_len)
        rl = np.where(op == OP_D, ln, 0)
        ok = (p >= s[rid] + 1) & (p + rl <= e[rid] - 1)
        rid, p, op, ln, rl = rid[ok], p[ok], op[ok], ln[ok], rl[ok]
        order = np.argsort(rid.astype(np.uint64) << np.uint64(34) | p.astype(np.uint64), kind="stable")
        rid, p, op, ln, rl = rid[order], p[order], op[order], ln[order], rl[order]
        # drop events that start before the previous event (same read) has ended + 1
        while True:
            same = np.concatenate(([False], rid[1:] == rid[:-1]))
            prev_end = np.concatenate(([0], (p + rl)[:-1]))
            bad = same & (p < prev_end + 1)
            if not bad.any():
                break
            keep = ~bad
            rid, p, op, ln, rl = rid[keep], p[keep], op[keep], ln[keep], rl[keep]
        same = np.concatenate(([False], rid[1:] == rid[:-1]))
        prev_end = np.where(same, np.concatenate(([0], (p + rl)[:-1])), s[rid])
        gap = p - prev_end                                   # M run before each event, >= 1
        nev = np.bincount(rid, minlength=nr)
        ev_off = np.cumsum(nev) - nev
        last_end = s.copy()
        has = nev > 0
        last_idx = (ev_off + nev - 1)[has]
        last_end[has] = (p + rl)[last_idx]
        tail_m = e - last_end                                # final M run, >= 1
        nw = 2 * nev + 1 + (lead > 0) + (trail > 0)
        woff = np.cumsum(nw) - nw
        words = np.zeros(int(nw.sum()), dtype=np.uint32)
        li = lead > 0
        words[woff[li]] = (lead[li] << 4 | OP_S).astype(np.uint32)
        base = woff + li                                     # first non-clip word of each read
        j = np.arange(rid.shape[0]) - ev_off[rid]            # event ordinal within its read
        words[base[rid] + 2 * j] = (gap << 4 | OP_M).astype(np.uint32)
        words[base[rid] + 2 * j + 1] = (ln << 4 | op).astype(np.uint32)
        words[base + 2 * nev] = (tail_m << 4 | OP_M).astype(np.uint32)
        ti = trail > 0
        words[(base + 2 * nev + 1)[ti]] = (trail[ti] << 4 | OP_S).astype(np.uint32)
        # split reads of kind B moved their start: re-sort the chunk by pos later (global sort below)
        out_pos.append(s); out_n.append(nw); out_words.append(words)
    pos = np.concatenate(out_pos) if out_pos else np.zeros(0, dtype=np.int64)
    ncig = np.concatenate(out_n) if out_n else np.zeros(0, dtype=np.int64)
    words = np.concatenate(out_words) if out_words else np.zeros(0, dtype=np.uint32)
    # coordinate sort (only B-type split reads are out of place); permute CIGAR blocks accordingly
    order = np.argsort(pos, kind="stable")
    if not np.array_equal(order, np.arange(pos.shape[0])):
        off = np.cumsum(ncig) - ncig
        _, widx = _expand_ranges(off[order], (off + ncig)[order])
        words = words[widx]
        pos, ncig = pos[order], ncig[order]
    return pos, ncig, words


def simulate(seed, contig_lens, coverage, read_len, small_rate, n_sv, *, sv_kind="alt", min_len=60, max_len=2000,
             min_gap=1500, log_uniform=False, carrier_frac=0.8, split_frac=0.05, margin=25000):
    """Whole read set + planted truth. SVs are spread over contigs proportionally to length."""
    rng = np.random.default_rng(seed)
    total = float(sum(contig_lens))
    pos_l, n_l, w_l, tid_off = [], [], [], [0]
    t_tid, t_pos, t_len, t_del = [], [], [], []
    left = n_sv
    for tid, L in enumerate(contig_lens):
        k = int(round(n_sv * L / total)) if tid < len(contig_lens) - 1 else left
        k = max(0, min(k, left)); left -= k
        sp, sl, sd = plant_svs(rng, L, k, min_len=min_len, max_len=max_len, min_gap=min_gap, margin=margin,
                               kind=sv_kind, log_uniform=log_uniform)
        nreads = int(L * coverage / read_len)
        pos, ncig, words = gen_contig(rng, L, nreads, read_len, small_rate, sp, sl, sd,
                                      carrier_frac=carrier_frac, split_frac=split_frac)
        pos_l.append(pos); n_l.append(ncig); w_l.append(words); tid_off.append(tid_off[-1] + pos.shape[0])
        t_tid.append(np.full(k, tid, dtype=np.int32)); t_pos.append(sp); t_len.append(sl); t_del.append(sd)
    pos = np.concatenate(pos_l).astype(np.uint32)
    ncig = np.concatenate(n_l).astype(np.uint32)
    words = np.concatenate(w_l)
    cigar_off = np.zeros(pos.shape[0] + 1, dtype=np.uint64)
    np.cumsum(ncig.astype(np.uint64), out=cigar_off[1:])
    rs = ReadSet(np.array(tid_off, dtype=np.int64), pos, None, ncig, cigar_off, words, list(contig_lens))
    rs.end = endpos_from_cigar(pos, cigar_off, words)
    truth = SVTruth(np.concatenate(t_tid), np.concatenate(t_pos), np.concatenate(t_len), np.concatenate(t_del))
    return rs, truth


# ----------------------------------------------------------------------------------------------------------
# BASELINE.json configs.  `scale` shrinks contig length / SV count for bounded samples.
CONFIGS = {
    # name: (contig_lens, coverage, read_len, small_rate, n_sv, kwargs)
    "c1": dict(contig_lens=[1_000_000], coverage=30, read_len=15000, small_rate=0.005, n_sv=500,
               min_gap=1500),
    "c2": dict(contig_lens=[64_000_000], coverage=30, read_len=15000, small_rate=0.005, n_sv=10_000,
               min_gap=1500),
    "c3": dict(contig_lens=[248_000_000, 242_000_000, 198_000_000, 190_000_000, 181_000_000, 171_000_000,
                            159_000_000, 145_000_000, 138_000_000, 134_000_000, 135_000_000, 133_000_000,
                            114_000_000, 107_000_000, 102_000_000, 90_000_000, 83_000_000, 80_000_000,
                            59_000_000, 64_000_000, 47_000_000, 51_000_000, 156_000_000, 73_000_000],
               coverage=30, read_len=15000, small_rate=0.005, n_sv=25_000, min_gap=1500),
    "c4": dict(contig_lens=[64_000_000], coverage=60, read_len=15000, small_rate=0.08, n_sv=20_000,
               sv_kind="ins", min_len=50, max_len=10_000, log_uniform=True, min_gap=1500),
    "c5": dict(contig_lens=[8_000_000], coverage=200, read_len=15000, small_rate=0.005, n_sv=2000,
               min_gap=50, min_len=60, max_len=300, margin=25000),
}


def make_config(name, seed=None, scale=1.0):
    cfg = dict(CONFIGS[name])
    if scale != 1.0:
        cfg["contig_lens"] = [max(200_000, int(L * scale)) for L in cfg["contig_lens"]]
        cfg["n_sv"] = max(1, int(cfg["n_sv"] * scale))
    if seed is None:
        seed = 20261104 + sorted(CONFIGS).index(name)
    if name == "c5":
        return _make_c5(seed, cfg)
    return simulate(seed, **cfg)


def _make_c5(seed, cfg):
    """Stress: tandem clusters -- 50 loci x 40 SVs spaced 50-300 bp => heavily overlapping windows."""
    rng = np.random.default_rng(seed)
    L = cfg["contig_lens"][0]
    n_loci = max(1, cfg["n_sv"] // 40)
    loci = np.sort(rng.integers(100_000, L - 100_000, n_loci))
    pos0, length, is_del = [], [], []
    for c in loci:
        p = c
        for j in range(40):
            ln = int(rng.integers(cfg["min_len"], cfg["max_len"] + 1))
            d = (j % 2 == 0)
            pos0.append(p); length.append(ln); is_del.append(d)
            p += (ln if d else 0) + int(rng.integers(50, 301))
    pos0 = np.array(pos0, dtype=np.int64); length = np.array(length, dtype=np.int64); is_del = np.array(is_del)
    order = np.argsort(pos0, kind="stable")
    pos0, length, is_del = pos0[order], length[order], is_del[order]
    nreads = int(L * cfg["coverage"] / cfg["read_len"])
    pos, ncig, words = gen_contig(rng, L, nreads, cfg["read_len"], cfg["small_rate"], pos0, length, is_del,
                                  carrier_frac=0.3, split_frac=0.05)
    cigar_off = np.zeros(pos.shape[0] + 1, dtype=np.uint64)
    np.cumsum(ncig.astype(np.uint64), out=cigar_off[1:])
    rs = ReadSet(np.array([0, pos.shape[0]], dtype=np.int64), pos.astype(np.uint32), None, ncig.astype(np.uint32),
                 cigar_off, words, [L])
    rs.end = endpos_from_cigar(rs.pos, cigar_off, words)
    return rs, SVTruth(np.zeros(pos0.shape[0], dtype=np.int32), pos0, length, is_del)


# ----------------------------------------------------------------------------------------------------------
def vcf_lines(truth: SVTruth, seed=1, *, end_first=False, pos_noise=40, chr_prefix="chr"):
    """VCF data lines shaped like utils/simvcf.py output (CIPOS;CIEND;END;SVELDT;SVTYPE appended to INFO).

    POS is the 1-based anchor of the planted breakpoint plus uniform noise in [-pos_noise, pos_noise]
    (an "imprecise" call).  end_first=True writes END= at the head of INFO so that the reference's
    strstr("END=") does not hit CIEND= first."""
    rng = np.random.default_rng(seed)
    lines = []
    noise = rng.integers(-pos_noise, pos_noise + 1, truth.pos0.shape[0])
    noise_e = rng.integers(-pos_noise, pos_noise + 1, truth.pos0.shape[0])
    ci = rng.integers(25, 150, (truth.pos0.shape[0], 4))
    for i in range(truth.pos0.shape[0]):
        pos = int(truth.pos0[i] + noise[i])
        ln = int(truth.length[i])
        if truth.is_del[i]:
            end = int(truth.pos0[i] + ln + noise_e[i])
            tail = f"CIPOS=-{ci[i,0]},{ci[i,1]};CIEND=-{ci[i,2]},{ci[i,3]};END={end};SVELDT=SIMULATED;SVTYPE=DEL"
            info = f"SVLEN=-{ln};{tail}"
            if end_first:
                info = f"END={end};SVLEN=-{ln};CIPOS=-{ci[i,0]},{ci[i,1]};CIEND=-{ci[i,2]},{ci[i,3]};SVELDT=SIMULATED;SVTYPE=DEL"
        else:
            end = pos + 1
            info = f"SVLEN={ln};CIPOS=-{ci[i,0]},{ci[i,1]};END={end};SVELDT=SIMULATED;SVTYPE=INS"
            if end_first:
                info = f"END={end};SVLEN={ln};CIPOS=-{ci[i,0]},{ci[i,1]};SVELDT=SIMULATED;SVTYPE=INS"
        lines.append(f"{chr_prefix}{int(truth.tid[i]) + 1}\t{pos}\tGoldStandard{i + 1}\tN\tN\t.\tPASS\t{info}\tGT\t0/1")
    return lines


VCF_HEADER = ("##fileformat=VCFv4.2\n"
              '##INFO=<ID=SVTYPE,Number=1,Type=String,Description="Type of structural variant">\n'
              "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tSAMPLE\n")


def write_vcf(path, lines, pad=0):
    """pad > 0 appends junk lines: the reference drops up to 2t queued lines at EOF."""
    with open(path, "w") as f:
        f.write(VCF_HEADER)
        for ln in lines:
            f.write(ln + "\n")
        for _ in range(pad):
            f.write("pad\n")


def cigar_string(words) -> str:
    return "".join(f"{int(w) >> 4}{OPCHARS[int(w) & 15]}" for w in words)


def write_sam(path, rs: ReadSet, seed=7, flags=None, chr_prefix="chr", random_seq=True):
    """SAM text for small sets (reference runs). SEQ/QUAL are uniform random so BGZF compresses realistically."""
    rng = np.random.default_rng(seed)
    bases = np.frombuffer(b"ACGT", dtype=np.uint8)
    with open(path, "wb") as f:
        f.write(b"@HD\tVN:1.6\tSO:coordinate\n")
        for t, L in enumerate(rs.contig_len):
            f.write(f"@SQ\tSN:{chr_prefix}{t + 1}\tLN:{L}\n".encode())
        for t in range(rs.n_tid):
            for r in range(int(rs.tid_off[t]), int(rs.tid_off[t + 1])):
                w = rs.cigar[int(rs.cigar_off[r]):int(rs.cigar_off[r + 1])]
                op = w & 15
                qlen = int(((w >> 4)[(op == OP_M) | (op == OP_I) | (op == OP_S) | (op == OP_EQ) | (op == OP_X)]).sum())
                if random_seq and qlen > 0:
                    seq = bases[rng.integers(0, 4, qlen)].tobytes()
                    qual = (rng.integers(10, 41, qlen) + 33).astype(np.uint8).tobytes()
                else:
                    seq, qual = b"*", b"*"
                flag = 0 if flags is None else int(flags[r])
                f.write(f"r{r}\t{flag}\t{chr_prefix}{t + 1}\t{int(rs.pos[r]) + 1}\t60\t{cigar_string(w)}\t*\t0\t0\t".encode())
                f.write(seq + b"\t" + qual + b"\n")


def windows_from_truth(truth: SVTruth, seed=1, *, pos_noise=40, wider=20000, median=10000, narrow=2000):
    """The refine_* windows the reference builds for vcf_lines(truth, seed, end_first=True) (audit.c:178,191-193),
    as SoA arrays: INS -> one SVTK_INS window, DEL -> SVTK_DEL_START + SVTK_DEL_END.  Same noise draws as vcf_lines.
    Returns dict(chrom,start,end,pos,kind) plus `record` (index of the SV each window belongs to)."""
    rng = np.random.default_rng(seed)
    n = truth.pos0.shape[0]
    noise = rng.integers(-pos_noise, pos_noise + 1, n)
    noise_e = rng.integers(-pos_noise, pos_noise + 1, n)
    pos = (truth.pos0 + noise).astype(np.int64)
    end = np.where(truth.is_del, truth.pos0 + truth.length + noise_e, pos + 1).astype(np.int64)
    chrom = truth.tid.astype(np.int64) + 1
    is_del = truth.is_del & (((end - pos) & 0xFFFFFFFF) > 50)      # uint32 arithmetic like audit.c:168,190: a noisy END < POS still prints
    nwin = np.where(truth.is_del, np.where(is_del, 2, 0), 1)
    first = np.cumsum(nwin) - nwin
    tot = int(nwin.sum())
    w_chrom = np.zeros(tot, np.int32); w_start = np.zeros(tot, np.int64); w_end = np.zeros(tot, np.int64)
    w_pos = np.zeros(tot, np.int64); w_kind = np.zeros(tot, np.uint8); w_rec = np.zeros(tot, np.int64)
    ins = ~truth.is_del
    i = first[ins]
    w_chrom[i] = chrom[ins]; w_start[i] = pos[ins] - median; w_end[i] = pos[ins] + median; w_pos[i] = pos[ins]; w_kind[i] = 2; w_rec[i] = np.flatnonzero(ins)
    d = first[is_del]
    w_chrom[d] = chrom[is_del]; w_start[d] = pos[is_del] - wider; w_end[d] = pos[is_del] + narrow; w_pos[d] = pos[is_del]; w_kind[d] = 0
    w_rec[d] = np.flatnonzero(is_del)
    w_chrom[d + 1] = chrom[is_del]; w_start[d + 1] = end[is_del] - narrow; w_end[d + 1] = end[is_del] + narrow; w_pos[d + 1] = end[is_del]
    w_kind[d + 1] = 1; w_rec[d + 1] = np.flatnonzero(is_del)
    m = 0xFFFFFFFF
    return dict(chrom=w_chrom, start=(w_start & m).astype(np.uint32), end=(w_end & m).astype(np.uint32),
                pos=(w_pos & m).astype(np.uint32), kind=w_kind, record=w_rec)


# ------------------------------------------------------------------------------------------------------------
# BAM + BAI writer (SAMv1 4.2, 5.2) for the file-level tests and bench legs: same records write_sam describes, BGZF blocks
# cut the way HTSlib cuts them (a record that does not fit starts a new block, bgzf.c bgzf_flush_try), zlib level 6.
def _reg2bin(beg, end):
    end -= 1
    for shift, base in ((14, 4681), (17, 585), (20, 73), (23, 9), (26, 1)):
        if beg >> shift == end >> shift:
            return base + (beg >> shift)
    return 0


def _bgzf_block(data):
    import struct
    import zlib
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    comp = c.compress(data) + c.flush()
    return (b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", len(comp) + 25) + comp +
            struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data)))


def write_bam(path, rs: ReadSet, seed=7, flags=None, chr_prefix="chr", truth=None, ins_err=0.02):
    """Coordinate-sorted BAM + `path`.bai of a ReadSet; SEQ / QUAL uniform random like write_sam.  With `truth`, the bases
    under an `I` op that carries a planted insertion (same contig, same length, cursor within 3 bases) are a copy of that
    insertion's own sequence with substitutions at rate ins_err, so that a consensus over the carriers means something.
    Returns the file size."""
    import struct
    rng = np.random.default_rng(seed)
    templates = {}
    if truth is not None:
        trng = np.random.default_rng(seed + 1)
        for k in np.flatnonzero(~truth.is_del):
            templates.setdefault(int(truth.tid[k]), []).append((int(truth.pos0[k]), int(truth.length[k]), trng.integers(0, 4, int(truth.length[k])).astype(np.uint8)))
    BLOCK = 0xff00
    text = "@HD\tVN:1.6\tSO:coordinate\n" + "".join(f"@SQ\tSN:{chr_prefix}{t + 1}\tLN:{L}\n" for t, L in enumerate(rs.contig_len))
    hdr = b"BAM\1" + struct.pack("<i", len(text)) + text.encode() + struct.pack("<i", rs.n_tid)
    for t, L in enumerate(rs.contig_len):
        nm = f"{chr_prefix}{t + 1}".encode() + b"\0"
        hdr += struct.pack("<i", len(nm)) + nm + struct.pack("<i", int(L))
    bins = [dict() for _ in range(rs.n_tid)]
    lin = [dict() for _ in range(rs.n_tid)]
    with open(path, "wb") as f:
        state = {"coff": 0, "buf": bytearray()}

        def flush():
            if state["buf"]:
                blk = _bgzf_block(bytes(state["buf"]))
                f.write(blk)
                state["coff"] += len(blk)
                state["buf"] = bytearray()

        def put(data):                     # HTSlib: flush first if the record does not fit; oversized records span blocks
            if len(state["buf"]) + len(data) > BLOCK:
                flush()
            v0 = state["coff"] << 16 | len(state["buf"])
            while len(state["buf"]) + len(data) > BLOCK:
                k = BLOCK - len(state["buf"])
                state["buf"] += data[:k]
                data = data[k:]
                flush()
            state["buf"] += data
            return v0, state["coff"] << 16 | len(state["buf"])

        put(hdr)
        flush()
        for t in range(rs.n_tid):
            for r in range(int(rs.tid_off[t]), int(rs.tid_off[t + 1])):
                w = rs.cigar[int(rs.cigar_off[r]):int(rs.cigar_off[r + 1])].astype("<u4")
                op = w & 15
                ql = int(((w >> 4)[(op == OP_M) | (op == OP_I) | (op == OP_S) | (op == OP_EQ) | (op == OP_X)]).sum())
                nib = rng.integers(0, 4, ql + (ql & 1)).astype(np.uint8)
                if t in templates and bool(((op == OP_I) & ((w >> 4) >= 50)).any()):
                    cur, q = int(rs.pos[r]), 0
                    for cw in w:
                        o, ln = int(cw) & 15, int(cw) >> 4
                        if o == OP_I and ln >= 50:
                            for p0, L, tpl in templates[t]:
                                if L == ln and abs(p0 - cur) <= 3:
                                    copy = tpl.copy()
                                    m = rng.random(ln) < ins_err
                                    copy[m] = (copy[m] + rng.integers(1, 4, int(m.sum()))) & 3
                                    nib[q:q + ln] = copy
                                    break
                        if o not in (OP_I, OP_S):
                            cur += ln
                        if o in (OP_M, OP_I, OP_S, OP_EQ, OP_X):
                            q += ln
                nib = (1 << nib).astype(np.uint8)                      # A C G T = 1 2 4 8
                if ql & 1:
                    nib[-1] = 0
                seq = (nib[0::2] << 4 | nib[1::2]).astype(np.uint8).tobytes()
                qual = rng.integers(10, 41, ql).astype(np.uint8).tobytes()
                name = f"r{r}".encode() + b"\0"
                flag = 0 if flags is None else int(flags[r])
                pos, end = int(rs.pos[r]), int(rs.end[r])
                if len(w) > 65535:
                    raise ValueError("write_bam: CIGARs beyond 65535 operations need the CG tag (use the HTSlib-written golden file)")
                body = struct.pack("<iiBBHHHiiii", t, pos, len(name), 60, _reg2bin(pos, end), len(w), flag, ql, -1, -1, 0) + name + w.tobytes() + seq + qual
                v0, v1 = put(struct.pack("<i", len(body)) + body)
                b = bins[t].setdefault(_reg2bin(pos, end), [])
                if b and b[-1][1] == v0:
                    b[-1][1] = v1
                else:
                    b.append([v0, v1])
                for win in range(pos >> 14, ((end - 1) >> 14) + 1):
                    lin[t].setdefault(win, v0)
        flush()
        f.write(_bgzf_block(b""))                                         # EOF marker block
        size = state["coff"] + 28
    with open(path + ".bai", "wb") as f:
        f.write(b"BAI\1" + struct.pack("<i", rs.n_tid))
        for t in range(rs.n_tid):
            f.write(struct.pack("<i", len(bins[t])))
            for b in sorted(bins[t]):
                f.write(struct.pack("<Ii", b, len(bins[t][b])))
                for v0, v1 in bins[t][b]:
                    f.write(struct.pack("<QQ", v0, v1))
            n_intv = (max(lin[t]) + 1) if lin[t] else 0
            f.write(struct.pack("<i", n_intv))
            prev = 0
            for win in range(n_intv):
                prev = lin[t].get(win, prev)                              # empty windows repeat the previous offset, like HTSlib
                f.write(struct.pack("<Q", prev))
    return size


# ------------------------------------------------------------------------------------------------------------
# The same kind of file, written fast: record bytes and BGZF compression by tools/libbgzfpack.so (C, zlib, pthreads; an input
# generator, not product code), block cutting and the index with numpy.  bench.py writes the FULL config-2 BAM (128 k reads,
# ~2 GB) with it, identically on both arms.
_QUERY_OPS = np.array([1, 1, 0, 0, 1, 0, 0, 1, 1, 0, 0, 0, 0, 0, 0, 0], dtype=np.int64)      # M I S = X consume the query


def _reg2bin_vec(beg, end):
    end = end - 1
    out = np.zeros(beg.shape[0], dtype=np.int64)
    done = np.zeros(beg.shape[0], dtype=bool)
    for shift, base in ((14, 4681), (17, 585), (20, 73), (23, 9), (26, 1)):
        hit = ~done & ((beg >> shift) == (end >> shift))
        out[hit] = base + (beg[hit] >> shift)
        done |= hit
    return out


def _packer():
    import ctypes
    import os as _os
    lib = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "tools", "libbgzfpack.so")
    if not _os.path.exists(lib):
        return None
    try:
        L = ctypes.CDLL(lib)
    except OSError:
        return None
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    L.bgzfpack.argtypes = [vp, vp, i32, i32, i32, vp, i64, vp]
    L.bgzfpack.restype = i32
    L.bam_records.argtypes = [i64, i64, vp, vp, vp, vp, vp, vp, vp, vp, ctypes.c_uint64, i32, vp]
    L.bam_records.restype = i32
    return L


def write_bam_fast(path, rs: ReadSet, seed=7, chr_prefix="chr", truth=None, ins_err=0.02, threads=None, level=6, chunk_reads=8192,
                   header_contigs=None, tid_shift=0, want_meta=False):
    """A BAM + BAI of the read set like write_bam's (same record fields, HTSlib's BGZF cutting rule, same index layout), fast
    enough for the full configs: ~15 s for config 2 on 16 cores, most of it zlib level 6.  Read names are fixed width (r%09d);
    SEQ / QUAL are uniform random, carriers of a planted insertion hold a mutated copy of its sequence (`truth`).
    header_contigs / tid_shift: the file is one PART of a larger job -- its header lists all of header_contigs, its records carry
    reference id tid + tid_shift -- so that parts written by different processes can be concatenated block-wise into one BAM
    (assemble_bam_parts).  want_meta: return what that needs instead of the size.
    Falls back to write_bam when tools/libbgzfpack.so has not been built.  Returns the file size."""
    import struct
    import os as _os
    L = _packer()
    if L is None:
        return write_bam(path, rs, seed=seed, chr_prefix=chr_prefix, truth=truth, ins_err=ins_err)
    threads = threads or max(1, min(64, (_os.cpu_count() or 2)))
    n = rs.n_reads
    BLOCK = 0xff00
    NAME = 11                                                  # "r%09d\0"
    tid = np.ascontiguousarray(np.repeat(np.arange(rs.n_tid, dtype=np.int32), np.diff(rs.tid_off)) + np.int32(tid_shift))
    hdr_contigs = list(header_contigs) if header_contigs is not None else list(rs.contig_len)
    n_ref_all = len(hdr_contigs)
    cig_off = np.ascontiguousarray(rs.cigar_off.astype(np.int64))
    ncig = rs.n_cigar.astype(np.int64)
    if ncig.size and int(ncig.max()) > 65535:
        raise ValueError("write_bam_fast: CIGARs beyond 65535 operations need the CG tag")
    qcs = np.zeros(rs.cigar.shape[0] + 1, dtype=np.int64)
    np.cumsum(_QUERY_OPS[rs.cigar & 15] * (rs.cigar >> 4).astype(np.int64), out=qcs[1:])
    qlen = np.ascontiguousarray(qcs[cig_off[1:]] - qcs[cig_off[:-1]]) if n else np.zeros(0, np.int64)
    del qcs
    rec_len = 4 + 32 + NAME + 4 * ncig + (qlen + 1) // 2 + qlen
    bins = _reg2bin_vec(rs.pos.astype(np.int64), rs.end.astype(np.int64)) if n else np.zeros(0, np.int64)
    bins16 = np.ascontiguousarray(bins.astype(np.uint16))
    ncig32 = np.ascontiguousarray(rs.n_cigar.astype(np.uint32))
    pos32 = np.ascontiguousarray(rs.pos.astype(np.uint32))
    cigar = np.ascontiguousarray(rs.cigar.astype(np.uint32))
    templates = {}
    has_big_ins = np.zeros(n, dtype=bool)
    if truth is not None and n:
        trng = np.random.default_rng(seed + 1)
        for k in np.flatnonzero(~truth.is_del):
            templates.setdefault(int(truth.tid[k]), []).append((int(truth.pos0[k]), int(truth.length[k]), trng.integers(0, 4, int(truth.length[k])).astype(np.uint8)))
        big = ((rs.cigar & 15) == OP_I) & ((rs.cigar >> 4) >= 50)
        bcs = np.zeros(rs.cigar.shape[0] + 1, dtype=np.int64)
        np.cumsum(big, out=bcs[1:])
        has_big_ins = (bcs[cig_off[1:]] - bcs[cig_off[:-1]]) > 0

    text = "@HD\tVN:1.6\tSO:coordinate\n" + "".join(f"@SQ\tSN:{chr_prefix}{t + 1}\tLN:{L_}\n" for t, L_ in enumerate(hdr_contigs))
    hdr = b"BAM\1" + struct.pack("<i", len(text)) + text.encode() + struct.pack("<i", n_ref_all)
    for t, L_ in enumerate(hdr_contigs):
        nm = f"{chr_prefix}{t + 1}".encode() + b"\0"
        hdr += struct.pack("<i", len(nm)) + nm + struct.pack("<i", int(L_))

    # ---- pass 1 (sizes only): the block every record starts / ends in.  HTSlib: a record that does not fit the current block
    # starts a new one; a record larger than a block fills blocks to the brim (bgzf.c: bgzf_flush_try, bgzf_write)
    hdr_blocks = max(1, -(-len(hdr) // BLOCK))
    b, fill = hdr_blocks, 0                                    # the header is flushed: records start a fresh block
    bs_l, is_l, be_l, ie_l = [0] * n, [0] * n, [0] * n, [0] * n
    starts = {}                                                # block -> stream offset of its first byte
    off = 0
    for r, Lr in enumerate(rec_len.tolist()):
        if fill + Lr > BLOCK and fill > 0:
            b += 1; fill = 0
        if fill == 0:
            starts[b] = off
        bs_l[r] = b; is_l[r] = fill
        rem, o2 = Lr, off
        while fill + rem > BLOCK:
            take = BLOCK - fill
            rem -= take; o2 += take
            b += 1; fill = 0
            starts[b] = o2
        fill += rem
        be_l[r] = b; ie_l[r] = fill
        off += Lr
    n_blocks = (b + 1) if fill > 0 else b                       # blocks [hdr_blocks, n_blocks) hold records
    if n and fill == 0:
        starts.pop(b, None)
    blk_start = np.array(bs_l, dtype=np.int64); in_start = np.array(is_l, dtype=np.int64)
    blk_end = np.array(be_l, dtype=np.int64); in_end = np.array(ie_l, dtype=np.int64)
    rec_off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(rec_len, out=rec_off[1:])
    blk_stream0 = np.zeros(n_blocks + 1, dtype=np.int64)
    for k in range(hdr_blocks, n_blocks):
        blk_stream0[k] = starts[k]
    blk_stream0[n_blocks] = rec_off[n]

    csize = np.zeros(n_blocks + 1, dtype=np.int64)
    rng = np.random.default_rng(seed)
    STR = 65536 + 64

    def pack(f, data, boff, k0):
        nb_out = boff.shape[0] - 1
        outb = np.empty(nb_out * STR, dtype=np.uint8)
        osz = np.zeros(nb_out, dtype=np.int64)
        if L.bgzfpack(data.ctypes.data, boff.ctypes.data, nb_out, level, threads, outb.ctypes.data, STR, osz.ctypes.data) != 0:
            raise RuntimeError("bgzfpack failed")
        for i in range(nb_out):
            f.write(outb[i * STR:i * STR + int(osz[i])].data)
        csize[k0:k0 + nb_out] = osz

    with open(path, "wb") as f:
        hb = np.frombuffer(hdr, dtype=np.uint8).copy()
        hoff = np.array([min(len(hdr), i * BLOCK) for i in range(hdr_blocks + 1)], dtype=np.int64)
        pack(f, hb, hoff, 0)
        carry = np.zeros(0, dtype=np.uint8)                    # bytes of the stream not yet written (tail of the last chunk)
        next_blk = hdr_blocks
        for c0 in range(0, n, chunk_reads):
            c1 = min(n, c0 + chunk_reads)
            base = int(rec_off[c0])
            d_off = base - carry.size                          # stream offset of data[0]
            data = np.empty(carry.size + int(rec_off[c1]) - base, dtype=np.uint8)
            data[:carry.size] = carry
            ro = np.ascontiguousarray(rec_off[c0:c1] - d_off)
            L.bam_records(c1 - c0, c0, tid[c0:].ctypes.data, pos32[c0:].ctypes.data, bins16[c0:].ctypes.data, ncig32[c0:].ctypes.data,
                          cig_off[c0:].ctypes.data, cigar.ctypes.data, qlen[c0:].ctypes.data, ro.ctypes.data, seed, threads, data.ctypes.data)
            # planted insertions: carriers hold a mutated copy of the insertion's own sequence
            for r in np.flatnonzero(has_big_ins[c0:c1]).tolist():
                rr = c0 + r
                t = int(tid[rr]) - int(tid_shift)
                if t not in templates:
                    continue
                w = rs.cigar[int(cig_off[rr]):int(cig_off[rr + 1])]
                cur, q = int(rs.pos[rr]), 0
                s0 = int(ro[r]) + 36 + NAME + 4 * int(ncig[rr])
                for cw in w.tolist():
                    o, ln = cw & 15, cw >> 4
                    if o == OP_I and ln >= 50:
                        for p0, Lt, tpl in templates[t]:
                            if Lt == ln and abs(p0 - cur) <= 3:
                                copy = tpl.copy()
                                mm = rng.random(ln) < ins_err
                                copy[mm] = (copy[mm] + rng.integers(1, 4, int(mm.sum()))) & 3
                                qi = q + np.arange(ln)
                                byte = s0 + (qi >> 1)
                                hi = (qi & 1) == 0
                                nibv = (1 << copy).astype(np.uint8)
                                data[byte[hi]] = (data[byte[hi]] & 0x0F) | (nibv[hi] << 4)       # two passes: a byte may take both nibbles
                                data[byte[~hi]] = (data[byte[~hi]] & 0xF0) | nibv[~hi]
                                break
                    if o not in (OP_I, OP_S):
                        cur += ln
                    if o in (OP_M, OP_I, OP_S, OP_EQ, OP_X):
                        q += ln
            # every block that ends inside the bytes built so far goes out now
            d_end = int(rec_off[c1])
            last_blk = next_blk
            while last_blk < n_blocks and int(blk_stream0[last_blk + 1]) <= d_end:
                last_blk += 1
            if last_blk > next_blk:
                boff = np.ascontiguousarray(blk_stream0[next_blk:last_blk + 1] - d_off, dtype=np.int64)
                pack(f, data, boff, next_blk)
            done_to = int(blk_stream0[last_blk])
            carry = data[done_to - d_off:].copy()
            next_blk = last_blk
        assert next_blk == n_blocks and carry.size == 0, (next_blk, n_blocks, carry.size)
        e = np.zeros(1, dtype=np.uint8)
        pack_eof = np.zeros(2, dtype=np.int64)                 # one empty block: the EOF marker
        pack(f, e, pack_eof, n_blocks)
    coff = np.zeros(n_blocks + 2, dtype=np.int64)
    np.cumsum(csize[:n_blocks + 1], out=coff[1:])
    size = int(coff[n_blocks + 1])
    # ---- index (SAMv1 5.2)
    v0 = (coff[blk_start] << 16) | in_start
    v1 = (coff[blk_end] << 16) | in_end
    sections = {}
    for t in range(rs.n_tid):
        a, z = int(rs.tid_off[t]), int(rs.tid_off[t + 1])
        sections[t + int(tid_shift)] = bai_section(bins[a:z], v0[a:z], v1[a:z], rs.pos[a:z], rs.end[a:z])
    with open(path + ".bai", "wb") as f:
        f.write(b"BAI\1" + struct.pack("<i", n_ref_all))
        for t in range(n_ref_all):
            f.write(sections.get(t, struct.pack("<ii", 0, 0)))
    if want_meta:
        return dict(size=size, hdr_bytes=int(coff[hdr_blocks]), v0=v0, v1=v1, bins=bins, n_ref_all=n_ref_all)
    return size


def bai_section(bins, v0, v1, pos, end, shift=0):
    """One reference's part of a .bai (SAMv1 5.2): bins with their chunk lists, then the 16 kb linear index.  bins / v0 / v1 / pos /
    end: per record, in file order.  shift: added to the compressed-offset half of every virtual offset (a part that moved)."""
    import struct
    n = int(bins.shape[0])
    if n == 0:
        return struct.pack("<ii", 0, 0)
    s0 = v0 + (int(shift) << 16)
    s1 = v1 + (int(shift) << 16)
    order = np.argsort(bins, kind="stable")
    bo, so0, so1 = bins[order], s0[order], s1[order]
    brk = np.ones(n, dtype=bool)
    brk[1:] = (bo[1:] != bo[:-1]) | (so0[1:] != so1[:-1])
    cs = np.flatnonzero(brk)
    ce = np.concatenate([cs[1:], [n]]) - 1
    ch_bin, ch_v0, ch_v1 = bo[cs], so0[cs], so1[ce]
    ub, first = np.unique(ch_bin, return_index=True)
    cnt = np.diff(np.concatenate([first, [ch_bin.shape[0]]]))
    out = [struct.pack("<i", ub.shape[0])]
    for bb, fi, cn in zip(ub.tolist(), first.tolist(), cnt.tolist()):
        out.append(struct.pack("<Ii", bb, cn))
        pair = np.empty(2 * cn, dtype="<u8")
        pair[0::2] = ch_v0[fi:fi + cn]; pair[1::2] = ch_v1[fi:fi + cn]
        out.append(pair.tobytes())
    p = np.asarray(pos).astype(np.int64); e_ = np.asarray(end).astype(np.int64)
    wlo, whi = p >> 14, ((e_ - 1) >> 14) + 1
    n_intv = int(whi.max())
    lin = np.full(n_intv, np.iinfo(np.int64).max, dtype=np.int64)
    own, widx = _expand_ranges(wlo, whi)
    np.minimum.at(lin, widx, s0[own])
    empty = lin == np.iinfo(np.int64).max
    last_filled = np.where(empty, -1, np.arange(n_intv))          # empty windows repeat the previous offset, like HTSlib
    np.maximum.accumulate(last_filled, out=last_filled)
    lin = np.where(last_filled < 0, 0, lin[np.maximum(last_filled, 0)])
    out.append(struct.pack("<i", n_intv))
    out.append(lin.astype("<u8").tobytes())
    return b"".join(out)


def assemble_bam_parts(out_path, part_paths, hdr_bytes, sizes):
    """One BAM out of parts written with the same header (write_bam_fast(..., header_contigs=, tid_shift=)): the header blocks of
    part 0, every part's record blocks, one EOF block.  BGZF blocks are independent, so this is a byte-level concatenation.
    Returns [offset of part k's record blocks in the new file minus their offset in the part] -- the shift of its virtual offsets."""
    import os as _os
    shifts, at = [], hdr_bytes
    with open(out_path, "wb") as f:
        with open(part_paths[0], "rb") as p0:
            f.write(p0.read(hdr_bytes))
        for pth, sz in zip(part_paths, sizes):
            n = sz - hdr_bytes - 28                          # the part minus its header blocks and its EOF block
            shifts.append(at - hdr_bytes)
            with open(pth, "rb") as src:
                src.seek(hdr_bytes)
                left = n
                while left > 0:
                    buf = src.read(min(left, 64 << 20))
                    if not buf:
                        raise IOError("short part file " + pth)
                    f.write(buf); left -= len(buf)
            at += n
        with open(part_paths[0], "rb") as p0:
            p0.seek(sizes[0] - 28)
            f.write(p0.read(28))
    return shifts
