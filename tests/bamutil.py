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
