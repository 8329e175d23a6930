"""TEST INFRASTRUCTURE ONLY -- a plain-Python BAM reader (zlib + struct): the checker for svtk_load_bam.

Restates what the reference gets from HTSlib for every record: bgzf_read_block / inflate (bgzf.c:988, 746), bam_read1
(sam.c:754-795) including the CG-tag long-CIGAR fix-up (bam_tag2cigar, sam.c:645-690) and bam_endpos (sam.c:638-643).
tools/make_bam_golden.py checks this reader against HTSlib's own SAM rendering of the same file (oracle/_ref/test_view)."""
import struct
import zlib

import numpy as np

REF_CONSUMING = (0x3C1A7 >> (np.arange(16) << 1)) & 2      # BAM_CIGAR_TYPE bit 1 (htslib/sam.h:137-158)


def bgzf_blocks(buf):
    """[(file offset, payload offset, payload length, isize)] of every BGZF block."""
    out, off = [], 0
    while off < len(buf):
        assert buf[off:off + 4] == b"\x1f\x8b\x08\x04", off
        xlen = struct.unpack_from("<H", buf, off + 10)[0]
        x, bsize = 0, None
        while x < xlen:
            si1, si2, slen = struct.unpack_from("<BBH", buf, off + 12 + x)
            if si1 == 66 and si2 == 67 and slen == 2:
                bsize = struct.unpack_from("<H", buf, off + 12 + x + 4)[0] + 1
            x += 4 + slen
        isize = struct.unpack_from("<I", buf, off + bsize - 4)[0]
        out.append((off, off + 12 + xlen, bsize - xlen - 20, isize))
        off += bsize
    return out


def inflate_all(buf):
    parts = []
    for _, p, n, isize in bgzf_blocks(buf):
        d = zlib.decompress(buf[p:p + n], -15)
        assert len(d) == isize
        parts.append(d)
    return b"".join(parts)


def _aux_cg(aux, n_cigar):
    """bam_aux_get(b, "CG") as far as bam_tag2cigar uses it: the real CIGAR words or None."""
    p = 0
    size = {b"A": 1, b"c": 1, b"C": 1, b"s": 2, b"S": 2, b"i": 4, b"I": 4, b"f": 4, b"d": 8}
    while p + 3 <= len(aux):
        tag, ty = aux[p:p + 2], aux[p + 2:p + 3]
        p += 3
        if ty in size:
            sz = size[ty]
        elif ty in (b"Z", b"H"):
            sz = aux.index(b"\0", p) + 1 - p
        elif ty == b"B":
            sub, cnt = aux[p:p + 1], struct.unpack_from("<I", aux, p + 1)[0]
            if tag == b"CG":
                if sub in (b"I", b"i") and n_cigar <= cnt < (1 << 29):
                    return np.frombuffer(aux, "<u4", cnt, p + 5)
                return None
            sz = 5 + size[sub] * cnt
        else:
            raise ValueError("bad aux type")
        if tag == b"CG":
            return None
        p += sz
    return None


def read_bam(path):
    """Every record with refID >= 0, in file order, as the svtk_reads arrays + the reference dictionary."""
    data = inflate_all(open(path, "rb").read())
    assert data[:4] == b"BAM\1"
    l_text = struct.unpack_from("<i", data, 4)[0]
    p = 8 + l_text
    n_ref = struct.unpack_from("<i", data, p)[0]
    p += 4
    refs = []
    for _ in range(n_ref):
        ln = struct.unpack_from("<i", data, p)[0]
        name = data[p + 4:p + 4 + ln].rstrip(b"\0").decode()
        refs.append((name, struct.unpack_from("<i", data, p + 4 + ln)[0]))
        p += 8 + ln
    tid, pos, end, ncig, cig, seqs = [], [], [], [], [], []
    n_records = 0
    while p < len(data):
        bs, refid, ps, lrn, _mapq, _bin, nc, flag, l_seq = struct.unpack_from("<iiiBBHHHi", data, p)
        n_records += 1
        if refid >= 0:
            c0 = p + 36 + lrn
            words = np.frombuffer(data, "<u4", nc, c0)
            if nc > 0 and ps >= 0 and (int(words[0]) & 15) == 4 and (int(words[0]) >> 4) == l_seq:
                real = _aux_cg(data[c0 + 4 * nc + (l_seq + 1) // 2 + l_seq:p + 4 + bs], nc)
                if real is not None:
                    words = real
            rlen = int((words >> 4)[REF_CONSUMING[words & 15] != 0].sum()) if len(words) else 0
            if (flag & 4) or rlen == 0:
                rlen = 1
            packed = np.frombuffer(data, np.uint8, (l_seq + 1) // 2, c0 + 4 * nc)
            nib = np.empty(2 * len(packed), np.uint8)
            nib[0::2] = packed >> 4; nib[1::2] = packed & 15              # bam_seqi: high nibble first
            seqs.append(nib[:l_seq])
            tid.append(refid); pos.append(ps & 0xFFFFFFFF); end.append((ps + rlen) & 0xFFFFFFFF); ncig.append(len(words)); cig.append(words)
        p += 4 + bs
    assert p == len(data)
    tid = np.asarray(tid, np.int64)
    ncig = np.asarray(ncig, np.uint32)
    off = np.zeros(len(ncig), np.uint64)
    if len(ncig):
        off[1:] = np.cumsum(ncig[:-1].astype(np.uint64))
    return dict(refs=refs, n_records=n_records, tid=tid, seq=seqs, tid_off=np.searchsorted(tid, np.arange(n_ref + 1)).astype(np.int64),
                pos=np.asarray(pos, np.uint32), end=np.asarray(end, np.uint32), n_cigar=ncig, cigar_off=off,
                cigar=np.concatenate(cig).astype(np.uint32) if cig else np.zeros(0, np.uint32))


NIB2CODE = np.full(16, 4, np.uint8)
NIB2CODE[[1, 2, 4, 8]] = [0, 1, 2, 3]                                     # =ACMGRSVTWYHKDBN -> abPOA's A0 C1 G2 T3 other 4


def ins_sequences(rd, chrom, start, end, kind, refined, tol=5, min_len=50):
    """Restatement of svtk_ins_sequences on read_bam() output: per INS window with refined >= 0, the inserted bases of the
    `I >= min_len` ops refine_ins would push (selection hts.c:3913-3916; walk + stop rule refinement.c:299-318) within tol of the
    refined position.  Returns [(window, [(read, ref_pos, codes)])] for windows with at least one sequence."""
    out = []
    for w in range(len(chrom)):
        if int(kind[w]) != 2 or int(refined[w]) < 0:
            continue
        ws, we = int(start[w]) & 0xFFFFFFFF, int(end[w]) & 0xFFFFFFFF
        t = int(chrom[w]) - 1
        beg, endq = (ws - 1) & 0xFFFFFFFF, (we - 1) & 0xFFFFFFFF
        if t < 0 or t >= len(rd["tid_off"]) - 1 or endq < beg:
            continue
        got = []
        for r in range(int(rd["tid_off"][t]), int(rd["tid_off"][t + 1])):
            p = int(rd["pos"][r])
            if p >= endq:
                break
            if not int(rd["end"][r]) > beg:
                continue
            cur, q = p, 0
            o = int(rd["cigar_off"][r])
            for cw in rd["cigar"][o:o + int(rd["n_cigar"][r])]:
                op, ln = int(cw) & 15, int(cw) >> 4
                if op == 1 and ln >= min_len and abs(cur - int(refined[w])) <= tol and q + ln <= len(rd["seq"][r]):
                    got.append((r, cur, NIB2CODE[rd["seq"][r][q:q + ln]]))
                if op not in (1, 4):
                    cur += ln
                if op in (0, 1, 4, 7, 8):
                    q += ln
                if cur > we:
                    break
        if got:
            out.append((w, got))
    return out
