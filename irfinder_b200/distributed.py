"""Region-sharded multi-GPU run: one process and one irf_ctx per GPU, torch.distributed for the exchange.

SURVEY.md §8e.  A coordinate-sorted BAM is cut at BGZF block boundaries into `world` byte ranges of equal
compressed size; BGZF offsets order records by coordinate, so each range is a genomic region.  The cut points
come from the file's BAI index when one is given (`bai=`: block offsets of its linear index and bins) and from
a scan for BGZF block headers otherwise; either way record framing inside a shard is speculative + verified
(kernels.cu, S2) and the first record of shard r is cross-checked against the last record of shard r-1
(`frame handshake`), so the index only chooses where to cut, it is never trusted for record boundaries.

Exchange steps (all integer, so outputs are identical for every GPU count):
  1. halo:      records whose mate lies in another shard are trimmed to fixed part + name + CIGAR and
                all-gathered; every rank pairs the gathered set and counts the pairs whose key it owns.
  2. junctions: the three JunctionCount hash tables are exported as dense (key, counts) lists, all-gathered
                and re-inserted (slot positions differ between ranks, so they cannot be reduced in place).
  3. additive:  span-point, ROI, chromosome and read counters are all-reduced (SUM) in place on the library's
                device buffers (NCCL over NVLink).
  4. depth:     region-owned (§8e).  Every measured segment has one owning rank, chosen from the genomic position
                of each rank's first record; a rank lists its non-zero difference counters in foreign segments
                (reads reaching over a cut, cross-shard pairs: KBs to MBs) and the owners add them.  Nothing
                table-sized crosses NVLink.  (A shard that is not a genomic region - unsorted input - would list
                most of its counters; above FOREIGN_DENSE_LIMIT entries the table is all-reduced instead.)
  5. rows:      irf_finalize computes the per-intron statistics of the OWNED introns only; the row tables (all
                other rows zero) are summed as 64-bit integers and irf_finalize_rows brings them to the host.
Afterwards every rank holds the complete tables; rank 0 writes the files.
"""
import numpy as np

_BGZF_MAGIC = bytes([31, 139, 8, 4])
FOREIGN_DENSE_LIMIT = 1 << 27   # foreign depth entries (all ranks together) above which the depth table is all-reduced instead


def _is_block_start(buf, off):
    """True when a BGZF block header with a BC subfield starts at buf[off]; returns (ok, block_size)."""
    if off + 18 > len(buf) or bytes(buf[off:off + 4]) != _BGZF_MAGIC:
        return False, 0
    xlen = int(buf[off + 10]) | int(buf[off + 11]) << 8
    x = 0
    while x + 4 <= xlen and off + 12 + x + 6 <= len(buf):
        si1, si2 = int(buf[off + 12 + x]), int(buf[off + 13 + x])
        slen = int(buf[off + 14 + x]) | int(buf[off + 15 + x]) << 8
        if si1 == 66 and si2 == 67 and slen == 2:
            return True, (int(buf[off + 16 + x]) | int(buf[off + 17 + x]) << 8) + 1
        x += 4 + slen
    return False, 0


def find_block_start(buf, approx, hops=3):
    """First BGZF block start at or after `approx`, verified by following `hops` BSIZE links (SAM v1 §4.1)."""
    n = len(buf)
    off = int(approx)
    while off + 18 <= n:
        j = off
        view = bytes(buf[off:min(n, off + (1 << 17))])
        k = view.find(_BGZF_MAGIC)
        if k < 0:
            off += max(1, len(view) - 3)
            continue
        j = off + k
        ok, cur = True, j
        for _ in range(hops):
            good, bs = _is_block_start(buf, cur)
            if not good:
                ok = False
                break
            cur += bs
            if cur >= n:
                break
        if ok:
            return j
        off = j + 1
    return n


def shard_cuts(buf, world):
    """Byte offsets [0, c1, ..., len(buf)]: BGZF-block-aligned cuts that balance compressed bytes."""
    n = len(buf)
    cuts = [0]
    for r in range(1, world):
        c = find_block_start(buf, n * r // world)
        cuts.append(max(c, cuts[-1]))
    cuts.append(n)
    return cuts


def read_bai_block_offsets(bai_path):
    """Compressed file offsets of BGZF blocks in which some alignment starts, from a BAI index (SAM v1 §5.2): the upper 48 bits
    of every linear-index entry and of every bin chunk begin.  Sorted, unique.  These are the cut candidates of §8e's
    "host reads the BAI": each is a block start by construction, so no scan of the BGZF bytes is needed."""
    import struct
    raw = open(bai_path, "rb").read()
    if raw[:4] != b"BAI\1":
        raise ValueError(f"{bai_path}: not a BAI index (bad magic)")
    p = 4
    n_ref, = struct.unpack_from("<i", raw, p)
    p += 4
    offs = set()
    for _ in range(n_ref):
        n_bin, = struct.unpack_from("<i", raw, p)
        p += 4
        for _ in range(n_bin):
            bin_id, n_chunk = struct.unpack_from("<Ii", raw, p)
            p += 8
            if bin_id != 37450:                       # the pseudo-bin holds counts, not offsets
                ch = np.frombuffer(raw, dtype="<u8", count=2 * n_chunk, offset=p)
                offs.update(int(x) >> 16 for x in ch[0::2])
            p += 16 * n_chunk
        n_intv, = struct.unpack_from("<i", raw, p)
        p += 4
        io = np.frombuffer(raw, dtype="<u8", count=n_intv, offset=p)
        offs.update(int(x) >> 16 for x in np.unique(io) if x)
        p += 8 * n_intv
    return sorted(offs)


def shard_cuts_from_bai(bai_path, buf, world):
    """Like shard_cuts, with the candidates taken from the BAI instead of a scan for block headers; a candidate that is not a
    BGZF block start in `buf` (index of another file, truncated file) raises."""
    n = len(buf)
    cand = [c for c in read_bai_block_offsets(bai_path) if 0 < c < n]
    cuts = [0]
    for r in range(1, world):
        want = n * r // world
        i = int(np.searchsorted(np.asarray(cand, dtype=np.int64), want)) if cand else 0
        near = [c for c in cand[max(0, i - 1): i + 1] if c > cuts[-1]]        # the candidates either side of the ideal cut
        if not near:
            near = [c for c in cand[i:] if c > cuts[-1]][:1]                      # sparse index: the next candidate, if any is left
        c = min(near, key=lambda x: abs(x - want)) if near else n
        if c < n and not _is_block_start(buf, c)[0]:
            raise ValueError(f"{bai_path}: offset {c} is not a BGZF block start in this BAM (index of another file?)")
        cuts.append(max(c, cuts[-1]))
    cuts.append(n)
    return cuts


class _DevArray:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = dict(shape=(int(n),), typestr=typestr, data=(int(ptr), False), version=3)


def _wrap(ptr, n, typestr, device):
    """Zero-copy torch view of `n` elements at address `ptr` (device memory for a cuda device, host memory for cpu)."""
    import ctypes
    import torch
    tdt = {"<i4": torch.int32, "<i8": torch.int64, "|u1": torch.uint8}[typestr]
    if n == 0 or not ptr:
        return torch.zeros(0, dtype=tdt, device=device)
    if torch.device(device).type == "cpu":   # the gloo tests drive this module with host-resident fake tables
        ndt = {"<i4": np.int32, "<i8": np.int64, "|u1": np.uint8}[typestr]
        raw = (ctypes.c_uint8 * (int(n) * np.dtype(ndt).itemsize)).from_address(int(ptr))
        return torch.from_numpy(np.frombuffer(raw, dtype=ndt))
    return torch.as_tensor(_DevArray(ptr, n, typestr), device=device)


def _sync(device):
    import torch
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize()


def _all_gather_var(t, group, dist):
    """all_gather of 1-D tensors with per-rank lengths; returns (list of tensors, lengths)."""
    import torch
    world = dist.get_world_size(group)
    n = torch.tensor([t.numel()], dtype=torch.int64, device=t.device)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n, group=group)
    lens = [int(x.item()) for x in ns]
    m = max(max(lens), 1)
    pad = torch.zeros(m, dtype=t.dtype, device=t.device)
    pad[: t.numel()] = t
    outs = [torch.zeros_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    return [o[:l] for o, l in zip(outs, lens)], lens


class ShardPlan:
    """What one rank feeds its context: the BAM header prefix, its own BGZF byte range and some look-ahead bytes of the
    right neighbour (so a record straddling the shard end can be read).  `data` holds own + look-ahead contiguously."""

    def __init__(self, header, data, n_own, rank, world, cuts=None):
        self.header, self.data, self.n_own, self.rank, self.world, self.cuts = header, data, int(n_own), rank, world, cuts

    @staticmethod
    def from_buffer(buf, rank, world, lookahead=1 << 18, header_probe=1 << 22, bai=None):
        """Every rank sees the whole BAM (host memory or np.memmap); cuts balance compressed bytes.  With `bai` (path of the
        file's BAI index) the cut points come from the index, otherwise from a scan for BGZF block headers."""
        arr = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
        n = arr.size
        cuts = shard_cuts_from_bai(bai, arr, world) if bai else shard_cuts(arr, world)
        lo, hi = cuts[rank], cuts[rank + 1]
        hi_look = hi if rank == world - 1 else min(n, hi + lookahead)
        return ShardPlan(arr[: min(n, header_probe)], arr[lo:hi_look], hi - lo, rank, world, cuts)

    @staticmethod
    def from_parts(part, header, rank, world, dist, device, group=None, lookahead=1 << 18, pin=True):
        """Every rank holds only ITS byte range of the BAM (`part`, e.g. produced by `irf_synth --chunk-range`); the
        look-ahead bytes come from the right neighbour through one all_gather of `lookahead` bytes per rank."""
        import torch
        part = np.frombuffer(part, dtype=np.uint8) if not isinstance(part, np.ndarray) else part
        head = torch.zeros(lookahead, dtype=torch.uint8)
        k = min(lookahead, part.size)
        head[:k] = torch.from_numpy(np.ascontiguousarray(part[:k]))
        meta = torch.tensor([k], dtype=torch.int64)
        dev = torch.device(device)
        heads = [torch.zeros(lookahead, dtype=torch.uint8, device=dev) for _ in range(world)]
        ks = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(heads, head.to(dev), group=group)
        dist.all_gather(ks, meta.to(dev), group=group)
        extra = heads[rank + 1][: int(ks[rank + 1].item())].cpu().numpy() if rank + 1 < world else np.zeros(0, np.uint8)
        base = part.base if isinstance(part.base, np.ndarray) else None
        if base is not None and base.ndim == 1 and base.dtype == np.uint8 and part.ctypes.data == base.ctypes.data and base.size >= part.size + extra.size:
            buf = base   # the caller left room behind its bytes: append the look-ahead in place (no second multi-GB pinned buffer)
        else:
            buf = torch.empty(part.size + extra.size + 64, dtype=torch.uint8, pin_memory=pin and torch.cuda.is_available()).numpy()
            buf[: part.size] = part
        buf[part.size: part.size + extra.size] = extra
        hdr = part if rank == 0 else (np.frombuffer(header, dtype=np.uint8) if not isinstance(header, np.ndarray) else header)
        return ShardPlan(hdr[: 1 << 26], buf[: part.size + extra.size], part.size, rank, world)


def run_plan(core, plan, device=None, group=None, restage=True):
    """Counts one ShardPlan per rank and merges; the merged, finalized tables are in `core` on every rank afterwards.

    restage=True streams the shard from host memory beside the inflate (end-to-end use).  restage=False copies it to the GPU
    the first time only and re-counts the resident bytes afterwards (what bench.py times as `value`)."""
    import torch
    import torch.distributed as dist
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    rank, world = plan.rank, plan.world
    if getattr(core, "_plan_header_id", None) != id(plan):   # BAM2blocks::readBamHeader once per (context, plan)
        hdr = np.ascontiguousarray(plan.header)
        probe = min(hdr.size, 1 << 20)
        while True:  # the header can span many blocks; grow the probe until it parses
            try:
                core.read_header(hdr.ctypes.data, probe)
                break
            except Exception:
                if probe >= hdr.size:
                    raise
                probe = min(hdr.size, probe * 4)
        core._plan_header_id = id(plan)
    data = plan.data if plan.data.flags["C_CONTIGUOUS"] else np.ascontiguousarray(plan.data)
    plan.data = data
    core.reset()
    if restage or getattr(core, "_plan_staged_id", None) != id(plan):
        if hasattr(core, "set_option") and dev.type == "cuda":
            core.set_option("stage_lazy", 1 if restage else 0)   # lazy: streamed to the GPU beside the inflate instead of copied first
        core.stage(data.ctypes.data, plan.n_own, data.size, has_header=(rank == 0))
        core._plan_staged_id = None if restage else id(plan)
    import time as _time
    _t = [_time.perf_counter()]
    phases = {}

    def _mark(name):
        _sync(dev)
        now = _time.perf_counter()
        phases[name] = round((now - _t[0]) * 1e3, 2)
        _t[0] = now
    core.count()
    _mark("count")
    info = dict(cuts=plan.cuts, recount=0, phase_ms=phases)
    if world > 1:
        # ---- frame handshake: my first record must be where my left neighbour's last record ends.  A recount changes the
        # exit a rank reports, so the values are exchanged again until nobody has to move (one round in practice).
        for _ in range(world + 1):
            e0, ex = core.frame_info()
            mine = torch.tensor([e0, ex], dtype=torch.int64, device=dev)
            allv = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allv, mine, group=group)
            bad = [r for r in range(1, world) if int(allv[r][0]) != int(allv[r - 1][1])]
            if not bad:
                break
            if rank in bad:
                core.set_option("entry0", int(allv[rank - 1][1]))
                core.reset()
                core.count()
                info["recount"] += 1
        else:
            raise RuntimeError("record framing of neighbouring shards does not agree")
        _mark("handshake")
        # ---- 1. halo
        hb, hnb, ho, hnr = core.export_halo()
        _sync(dev)
        parts, blens = _all_gather_var(_wrap(hb, hnb, "|u1", dev), group, dist)
        offs, olens = _all_gather_var(_wrap(ho, hnr, "<i8", dev), group, dist)
        starts = np.concatenate([[0], np.cumsum(blens)[:-1]]).astype(np.int64)
        all_bytes = torch.cat(parts + [torch.zeros(64, dtype=torch.uint8, device=dev)])
        all_offs = torch.cat([o + int(s) for o, s in zip(offs, starts)]) if sum(olens) else torch.zeros(0, dtype=torch.int64, device=dev)
        _sync(dev)
        core.import_halo(all_bytes.data_ptr(), int(sum(blens)), all_offs.data_ptr() if all_offs.numel() else 0, int(all_offs.numel()), rank, world)
        info["halo_records"] = int(all_offs.numel())
        _mark("halo")
        # ---- 2. junction tables: the three exports travel in ONE gather ([n0, n1, n2 | keys | counts as 64-bit words])
        exp = [core.export_junctions(which) for which in range(3)]
        _sync(dev)
        packed = torch.cat([torch.tensor([e[2] for e in exp], dtype=torch.int64, device=dev)]
                           + [_wrap(e[0], e[2], "<i8", dev) for e in exp] + [_wrap(e[1], e[2], "<i8", dev) for e in exp])
        got, _ = _all_gather_var(packed, group, dist)
        _sync(dev)
        for which in range(3):
            ks, cs = [], []
            for r in range(world):
                if r == rank:
                    continue
                n = [int(x) for x in got[r][:3].tolist()]
                k0 = 3 + sum(n[:which])
                c0 = 3 + sum(n) + sum(n[:which])
                ks.append(got[r][k0:k0 + n[which]])
                cs.append(got[r][c0:c0 + n[which]])
            k, c = torch.cat(ks).contiguous(), torch.cat(cs).contiguous()
            if k.numel():
                core.import_junctions(which, k.data_ptr(), c.data_ptr(), int(k.numel()))
        _mark("junctions")
        # ---- 3. small additive tables, in place on the library's buffers
        for name, typestr in (("spans", "<i4"), ("roi", "<i4"), ("chr", "<i4"), ("stats", "<i8")):
            p, cnt, _ = core.device_table(name)
            if cnt:
                dist.all_reduce(_wrap(p, cnt, typestr, dev), op=dist.ReduceOp.SUM, group=group)
        _mark("small_tables")
        # ---- 4. depth: region-owned
        firsts = getattr(plan, "_firsts", None)          # the first record of every shard: a property of the plan's bytes, gathered once
        if firsts is None or info["recount"]:
            rid, pos = core.first_record()
            mine = torch.tensor([rid, pos], dtype=torch.int64, device=dev)
            got = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(got, mine, group=group)
            firsts = plan._firsts = [(int(f[0]), int(f[1])) for f in got]
        core.owner_plan([f[0] for f in firsts], [f[1] for f in firsts], rank)
        ip, vp, nf = core.export_foreign_depth()
        nfs = torch.tensor([nf], dtype=torch.int64, device=dev)
        dist.all_reduce(nfs, op=dist.ReduceOp.SUM, group=group)
        info["foreign_depth_entries"] = int(nfs.item())
        if info["foreign_depth_entries"] > FOREIGN_DENSE_LIMIT:     # not region-shaped shards: the dense merge is cheaper
            p, cnt, _ = core.device_table("diff")
            dist.all_reduce(_wrap(p, cnt, "<i4", dev), op=dist.ReduceOp.SUM, group=group)
        elif info["foreign_depth_entries"]:                         # (nothing to exchange when no counter crossed an ownership boundary)
            packed = torch.cat([_wrap(ip, nf, "<i8", dev), _wrap(vp, nf, "<i4", dev).to(torch.int64)])   # one gather: [indices | values]
            got, _ = _all_gather_var(packed, group, dist)
            ai = torch.cat([g[: g.numel() // 2] for g in got]).contiguous()
            av = torch.cat([g[g.numel() // 2:] for g in got]).to(torch.int32).contiguous()
            _sync(dev)
            if ai.numel():
                core.import_depth(ai.data_ptr(), av.data_ptr(), int(ai.numel()))
        _mark("depth_exchange")
    core.finalize()
    _mark("finalize")
    if world > 1:
        # ---- 5. rows of the owned introns -> every rank (integer SUM: exactly one rank holds a non-zero row)
        for name in ("rows_nondir",) + (("rows_dir",) if core.directionality != 0 else ()):
            p, cnt, _ = core.device_table(name)
            if cnt:
                dist.all_reduce(_wrap(p, cnt, "<i8", dev), op=dist.ReduceOp.SUM, group=group)
        _sync(dev)
        core.finalize_rows()
        _mark("rows")
    return info


def run_sharded(core, buf, rank, world, device=None, group=None, lookahead=1 << 18, bai=None):
    """Counts `buf` (the whole BGZF BAM in host memory, same bytes on every rank) across `world` ranks."""
    return run_plan(core, ShardPlan.from_buffer(buf, rank, world, lookahead, bai=bai), device, group)


def run_sharded_local(cores, buf, lookahead=1 << 18, header_probe=1 << 20, bai=None):
    """The same protocol with every "rank" a context in THIS process on the current GPU (no torch.distributed).

    Used by the single-GPU tests to check that a sharded run is byte-identical to an unsharded one, and usable to walk a
    BAM larger than device memory shard by shard.  `cores` are IRFinderCore objects with the same references loaded.
    """
    import torch
    world = len(cores)
    dev = torch.device("cuda", torch.cuda.current_device())
    arr = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    base, n = arr.ctypes.data, arr.size
    cuts = shard_cuts_from_bai(bai, arr, world) if bai else shard_cuts(arr, world)
    info = dict(cuts=cuts, recount=0, halo_records=0)
    frames = []
    for r, core in enumerate(cores):
        probe = header_probe
        while True:
            try:
                core.read_header(base, min(n, probe))
                break
            except Exception:
                if probe >= n:
                    raise
                probe *= 4
        lo, hi = cuts[r], cuts[r + 1]
        hi_look = hi if r == world - 1 else min(n, find_block_start(arr, min(n, hi + lookahead)))
        core.reset()
        core.stage(base + lo, hi - lo, hi_look - lo, has_header=(r == 0))
        core.count()
        e0, ex = core.frame_info()
        if r > 0 and e0 != frames[-1][1]:
            core.set_option("entry0", frames[-1][1])
            core.reset()
            core.count()
            e0, ex = core.frame_info()
            info["recount"] += 1
        frames.append((e0, ex))
    if world > 1:
        halos = [c.export_halo() for c in cores]
        torch.cuda.synchronize()
        parts = [_wrap(h[0], h[1], "|u1", dev).clone() for h in halos]
        offs, start = [], 0
        for h in halos:
            offs.append(_wrap(h[2], h[3], "<i8", dev).clone() + start)
            start += h[1]
        all_bytes = torch.cat(parts + [torch.zeros(64, dtype=torch.uint8, device=dev)])
        all_offs = torch.cat(offs)
        torch.cuda.synchronize()
        info["halo_records"] = int(all_offs.numel())
        for r, core in enumerate(cores):
            core.import_halo(all_bytes.data_ptr(), start, all_offs.data_ptr() if all_offs.numel() else 0, int(all_offs.numel()), r, world)
        for which in range(3):
            exp = []
            for c in cores:
                kp, cp, nk = c.export_junctions(which)
                torch.cuda.synchronize()
                exp.append((_wrap(kp, nk, "<i8", dev).clone(), _wrap(cp, nk * 2, "<i4", dev).clone(), nk))
            torch.cuda.synchronize()
            for r, core in enumerate(cores):
                for q, (k, c, nk) in enumerate(exp):
                    if q != r and nk:
                        core.import_junctions(which, k.data_ptr(), c.data_ptr(), nk)
        for name, typestr in (("spans", "<i4"), ("roi", "<i4"), ("chr", "<i4"), ("stats", "<i8")):
            views = [_wrap(*c.device_table(name)[:2], typestr, dev) for c in cores]
            if views[0].numel():
                total = torch.stack(views).sum(0)
                for v in views:
                    v.copy_(total)
        # depth: region-owned, as in run_plan
        firsts = [c.first_record() for c in cores]
        for r, core in enumerate(cores):
            core.owner_plan([f[0] for f in firsts], [f[1] for f in firsts], r)
        lists = []
        for c in cores:
            ip, vp, nf = c.export_foreign_depth()
            torch.cuda.synchronize()
            lists.append((_wrap(ip, nf, "<i8", dev).clone(), _wrap(vp, nf, "<i4", dev).clone()))
        ai, av = torch.cat([l[0] for l in lists]).contiguous(), torch.cat([l[1] for l in lists]).contiguous()
        info["foreign_depth_entries"] = int(ai.numel())
        torch.cuda.synchronize()
        if ai.numel():
            for core in cores:
                core.import_depth(ai.data_ptr(), av.data_ptr(), int(ai.numel()))
        torch.cuda.synchronize()
    for core in cores:
        core.finalize()
    if world > 1:
        for name in ("rows_nondir",) + (("rows_dir",) if cores[0].directionality != 0 else ()):
            views = [_wrap(*c.device_table(name)[:2], "<i8", dev) for c in cores]
            if views[0].numel():
                total = torch.stack(views).sum(0)
                for v in views:
                    v.copy_(total)
        torch.cuda.synchronize()
        for core in cores:
            core.finalize_rows()
    return info
