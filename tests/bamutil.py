"""Minimal BAM/BGZF writer + reader (stdlib only) for hand-made fixtures.  Formats: SAM v1 §4."""
import struct
import zlib

BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")
OPS = "MIDNSHP=X"


def cigar(s):
    """'40M1000N60M' -> [(40,'M'),(1000,'N'),(60,'M')]"""
    out, num = [], ""
    for ch in s:
        if ch.isdigit():
            num += ch
        else:
            out.append((int(num), ch))
            num = ""
    return out


def record(name, refid, pos, cig, flag, mrefid=-1, mpos=-1, tlen=0, lseq=None):
    ops = cigar(cig) if cig else []
    if lseq is None:
        lseq = sum(n for n, c in ops if c in "MIS=X")
    nm = name.encode() + b"\0"
    body = struct.pack("<iiBBHHHiiii", refid, pos, len(nm), 255, 4680, len(ops), flag, lseq, mrefid, mpos, tlen)
    body += nm + b"".join(struct.pack("<I", n << 4 | OPS.index(c)) for n, c in ops)
    body += bytes((i * 37 + 17) & 0xFF for i in range((lseq + 1) // 2)) + bytes(30 + (i % 11) for i in range(lseq))
    body += b"NHC\x01"
    return struct.pack("<i", len(body)) + body


def header(refs, sorted_=False):
    text = ("@HD\tVN:1.6\tSO:%s\n" % ("coordinate" if sorted_ else "unsorted")).encode()
    for n, l in refs:
        text += b"@SQ\tSN:" + n.encode() + b"\tLN:" + str(l).encode() + b"\n"
    h = b"BAM\1" + struct.pack("<i", len(text)) + text + struct.pack("<i", len(refs))
    for n, l in refs:
        h += struct.pack("<i", len(n) + 1) + n.encode() + b"\0" + struct.pack("<i", l)
    return h


def bgzf(data, block=65280, level=6):
    out = b""
    for i in range(0, len(data), block):
        chunk = data[i:i + block]
        c = zlib.compressobj(level, zlib.DEFLATED, -15)
        cd = c.compress(chunk) + c.flush()
        bs = len(cd) + 25
        out += bytes([31, 139, 8, 4, 0, 0, 0, 0, 0, 255, 6, 0, 66, 67, 2, 0, bs & 255, bs >> 8]) + cd + struct.pack("<II", zlib.crc32(chunk), len(chunk))
    return out


def write_bam(path, refs, records, sorted_=False, block=65280, level=6):
    raw = header(refs, sorted_) + b"".join(records)
    with open(path, "wb") as f:
        f.write(bgzf(raw, block, level) + BGZF_EOF)
    return raw


def inflate_file(path):
    data = open(path, "rb").read()
    out, d = b"", zlib.decompressobj(31)
    while data:
        out += d.decompress(data)
        data = d.unused_data
        if not data:
            break
        d = zlib.decompressobj(31)
    return out


def read_records(raw):
    """Yields (refid, pos, flag, name, cigar_string, mpos, raw_bytes) from inflated BAM bytes."""
    assert raw[:4] == b"BAM\1"
    l_text, = struct.unpack_from("<i", raw, 4)
    p = 8 + l_text
    n_ref, = struct.unpack_from("<i", raw, p)
    p += 4
    refs = []
    for _ in range(n_ref):
        ln, = struct.unpack_from("<i", raw, p)
        name = raw[p + 4:p + 4 + ln - 1].decode()
        lr, = struct.unpack_from("<i", raw, p + 4 + ln)
        refs.append((name, lr))
        p += 8 + ln
    recs = []
    while p < len(raw):
        bs, = struct.unpack_from("<i", raw, p)
        refid, pos, lname, mapq, bin_, ncig, flag, lseq, mref, mpos, tlen = struct.unpack_from("<iiBBHHHiiii", raw, p + 4)
        name = raw[p + 36:p + 36 + lname - 1].decode()
        cg = "".join("%d%s" % (v >> 4, OPS[v & 15]) for v in struct.unpack_from("<%dI" % ncig, raw, p + 36 + lname))
        recs.append((refid, pos, flag, name, cg, mpos, raw[p:p + 4 + bs]))
        p += 4 + bs
    return refs, recs


def reg2bin(beg, end):
    """SAM v1 §5.3."""
    end -= 1
    for shift, base in ((14, 4681), (17, 585), (20, 73), (23, 9), (26, 1)):
        if beg >> shift == end >> shift:
            return base + (beg >> shift)
    return 0


def write_bai(bam_path, bai_path=None):
    """Indexes a coordinate-sorted BAM the way `samtools index` does (bins with merged chunks + 16 kb linear index).
    Test infrastructure: walks the BGZF blocks with zlib on the host."""
    data = open(bam_path, "rb").read()
    blocks, raw, off = [], bytearray(), 0           # (compressed offset, inflated offset)
    while off + 18 <= len(data):
        bs = (data[off + 16] | data[off + 17] << 8) + 1
        blocks.append((off, len(raw)))
        raw += zlib.decompress(data[off + 18:off + bs - 8], -15)
        off += bs
    starts = [b[1] for b in blocks]

    def voff(u):
        import bisect
        i = bisect.bisect_right(starts, u) - 1
        while i + 1 < len(blocks) and blocks[i + 1][1] == u:   # a record starting exactly at a block end belongs to the next block
            i += 1
        return blocks[i][0] << 16 | (u - blocks[i][1])

    l_text, = struct.unpack_from("<i", raw, 4)
    p = 8 + l_text
    n_ref, = struct.unpack_from("<i", raw, p)
    p += 4
    for _ in range(n_ref):
        ln, = struct.unpack_from("<i", raw, p)
        p += 4 + ln + 4
    bins = [dict() for _ in range(n_ref)]
    lin = [dict() for _ in range(n_ref)]
    while p + 4 <= len(raw):
        bsz, = struct.unpack_from("<i", raw, p)
        refid, pos, l_name, _mq, _bin, n_cig, flag, l_seq = struct.unpack_from("<iiBBHHHi", raw, p + 4)
        end = pos
        for k in range(n_cig):
            v, = struct.unpack_from("<I", raw, p + 36 + l_name + 4 * k)
            if "MIDNSHP=X"[v & 15] in "MDN=X":
                end += v >> 4
        end = max(end, pos + 1)
        if refid >= 0:
            b, e = voff(p), voff(p + 4 + bsz)
            ch = bins[refid].setdefault(reg2bin(pos, end), [])
            if ch and ch[-1][1] >> 16 == b >> 16:
                ch[-1][1] = e
            else:
                ch.append([b, e])
            for w in range(pos >> 14, ((end - 1) >> 14) + 1):
                lin[refid].setdefault(w, b)
        p += 4 + bsz
    out = b"BAI\1" + struct.pack("<i", n_ref)
    for r in range(n_ref):
        out += struct.pack("<i", len(bins[r]))
        for b in sorted(bins[r]):
            out += struct.pack("<Ii", b, len(bins[r][b])) + b"".join(struct.pack("<QQ", x, y) for x, y in bins[r][b])
        n_intv = max(lin[r]) + 1 if lin[r] else 0
        out += struct.pack("<i", n_intv)
        last = 0
        for w in range(n_intv):
            last = lin[r].get(w, last)
            out += struct.pack("<Q", last)
    bai_path = bai_path or bam_path + ".bai"
    with open(bai_path, "wb") as f:
        f.write(out)
    return bai_path
