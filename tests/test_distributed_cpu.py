"""CPU suite (gloo, world_size 2): the shard planner and the multi-rank merge protocol of irfinder_b200.distributed,
driven with a host-resident fake context (the real context needs a GPU; its sharded run is covered by the -m gpu tests)."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_shard_cuts_land_on_bgzf_blocks(built, tmp_path):
    from irfinder_b200 import synth
    from irfinder_b200.distributed import find_block_start, shard_cuts
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=80, pairs=20000, readlen=100, seed=41, chunks=5, collated=False)
    buf = np.fromfile(meta["sorted"], dtype=np.uint8)
    starts, off = set(), 0
    while off < buf.size:            # ground truth by walking BSIZE links from offset 0
        starts.add(off)
        off += (int(buf[off + 16]) | int(buf[off + 17]) << 8) + 1
    for world in (1, 2, 3, 4, 8):
        cuts = shard_cuts(buf, world)
        assert cuts[0] == 0 and cuts[-1] == buf.size and cuts == sorted(cuts) and len(cuts) == world + 1
        assert all(c in starts or c == buf.size for c in cuts)
        sizes = np.diff(cuts)
        assert sizes.max() - sizes.min() < 3 * 65536 + buf.size // (world * 50)
    # a magic byte sequence inside compressed data must not be mistaken for a block start
    fake = np.concatenate([buf[:100000], np.frombuffer(bytes([31, 139, 8, 4, 0, 0, 0, 0, 0, 255, 6, 0, 66, 67, 2, 0, 99, 0]), dtype=np.uint8), buf[100000:]])
    assert find_block_start(fake, 100000) != 100000


class FakeCore:
    """Host-memory stand-in with the IRFinderCore surface run_sharded() uses.  Tables are deterministic functions of the rank."""
    def __init__(self, rank, world):
        self.rank, self.world, self.calls = rank, world, []
        self.t = {"diff": np.arange(40, dtype=np.int32) * (rank + 1), "spans": np.full(6, rank + 1, np.int32), "roi": np.zeros(0, np.int32),
                  "chr": np.array([rank, 1], np.int32), "stats": np.arange(9, dtype=np.int64) + rank}
        # junction tables: rank-specific keys plus one shared key
        self.j = [{1000 + which: [1, 0], 5000 + 10 * rank + which: [rank + 1, 2]} for which in range(3)]
        self.halo = (np.arange(8 * (rank + 1), dtype=np.uint8), np.array([0, 8][: rank + 1] if rank == 0 else [0, 8], np.int64))
        self.imported_halo = None
        self.keep = []
        self.directionality = 0
        self.rows = np.zeros(8, np.int64)
        self.t["rows_nondir"] = self.rows
        self.owner = None
    def read_header(self, p, n): self.calls.append("header")
    def reset(self): self.calls.append("reset")
    def stage(self, p, own, total, has_header=True): self.calls.append(("stage", own, total, has_header))
    def count(self): self.calls.append("count")
    def set_option(self, k, v): self.calls.append(("opt", k, v))
    def frame_info(self): return (7 if self.rank else 0), 7   # consistent hand-shake: my entry == left neighbour's exit
    def finalize(self):
        self.calls.append("finalize")
        for j in range(8):            # owner mode: only the rows of the introns this rank owns
            if j * self.world // 8 == self.rank:
                self.rows[j] = (j + 1) * 100
    def finalize_rows(self): self.calls.append("finalize_rows")
    # region-owned depth: slot i of the 40 belongs to rank i * world // 40
    def first_record(self): return self.rank, 1000 * self.rank
    def owner_plan(self, refids, positions, rank): self.owner = (list(refids), list(positions), rank)
    def _owner_of(self, i): return (i % 40) * self.world // 40
    def export_foreign_depth(self):
        d = self.t["diff"]
        idx = np.array([i for i in range(d.size) if d[i] != 0 and self._owner_of(i) != self.rank], np.int64)
        val = d[idx].astype(np.int32)
        self.keep += [idx, val]
        return (idx.ctypes.data if idx.size else 0), (val.ctypes.data if val.size else 0), idx.size
    def import_depth(self, ip, vp, n):
        import ctypes
        idx = np.frombuffer((ctypes.c_int64 * n).from_address(ip), np.int64); val = np.frombuffer((ctypes.c_int32 * n).from_address(vp), np.int32)
        for i, v in zip(idx, val):
            if self._owner_of(int(i)) == self.rank:
                self.t["diff"][int(i)] += int(v)
    def device_table(self, name):
        a = self.t[name]
        return (a.ctypes.data if a.size else 0), a.size, a.itemsize
    def export_junctions(self, which):
        k = np.array(sorted(self.j[which]), np.int64); c = np.array([self.j[which][x] for x in sorted(self.j[which])], np.int32).reshape(-1)
        self.keep += [k, c]
        return k.ctypes.data, c.ctypes.data, k.size
    def import_junctions(self, which, kp, cp, n):
        import ctypes
        k = np.frombuffer((ctypes.c_int64 * n).from_address(kp), np.int64); c = np.frombuffer((ctypes.c_int32 * (2 * n)).from_address(cp), np.int32).reshape(-1, 2)
        for key, cc in zip(k, c):
            e = self.j[which].setdefault(int(key), [0, 0]); e[0] += int(cc[0]); e[1] += int(cc[1])
    def export_halo(self):
        b, o = self.halo
        return b.ctypes.data, b.size, o.ctypes.data, o.size
    def import_halo(self, bp, nb, op, nr, owner, n_owners):
        import ctypes
        self.imported_halo = (bytes((ctypes.c_uint8 * nb).from_address(bp)), list(np.frombuffer((ctypes.c_int64 * nr).from_address(op), np.int64)) if nr else [], owner, n_owners)


def _worker(rank, world, port, path, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from irfinder_b200.distributed import run_sharded
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    buf = np.fromfile(path, dtype=np.uint8)
    core = FakeCore(rank, world)
    info = run_sharded(core, buf, rank, world, device="cpu")
    assert core.owner == ([0, 1], [0, 1000], rank)
    q.put((rank, {k: v.tolist() for k, v in core.t.items()}, core.j, core.imported_halo, core.calls, info["cuts"]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_merge_protocol_gloo(built, tmp_path):
    import torch.multiprocessing as mp
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=60, pairs=8000, readlen=100, seed=42, chunks=3, collated=False)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, meta["sorted"], q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted([q.get(timeout=120) for _ in ps])
    [p.join(timeout=60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    (r0, t0, j0, h0, c0, cuts0), (r1, t1, j1, h1, c1, cuts1) = res
    assert cuts0 == cuts1 and len(cuts0) == 3
    # small additive tables: both ranks end with the element-wise sum
    for name in ("spans", "chr", "stats", "rows_nondir"):
        assert t0[name] == t1[name], name
    assert t0["spans"] == [3] * 6 and t0["chr"] == [1, 2] and t0["stats"] == (2 * np.arange(9) + 1).tolist()
    # depth is region-owned: each rank holds the complete sum in the slots it owns (rank 0: 0..19, rank 1: 20..39) and only its own
    # contribution elsewhere; nothing table-sized was exchanged
    want = (np.arange(40) * 3).tolist()
    assert t0["diff"][:20] == want[:20] and t1["diff"][20:] == want[20:]
    assert t0["diff"][20:] == np.arange(20, 40).tolist() and t1["diff"][:20] == (np.arange(20) * 2).tolist()
    # rows: each rank computed its own introns, the integer sum gives both the whole table
    assert t0["rows_nondir"] == [(j + 1) * 100 for j in range(8)]
    # junction tables: union of keys, shared key summed, identical on both ranks
    assert j0 == j1
    assert j0[0][1000] == [2, 0] and j0[0][5000] == [1, 2] and j0[0][5010] == [2, 2]
    # halo: every rank receives all records, offsets rebased by the byte length of the lower ranks, and its own owner id
    assert h0[0] == h1[0] == bytes(range(8)) + bytes(range(16)) and h0[1] == h1[1] == [0, 0 + 8, 8 + 8][:1] + [8, 16]
    assert (h0[2], h0[3]) == (0, 2) and (h1[2], h1[3]) == (1, 2)
    # only rank 0 stages with the header; both finalize last
    assert ("stage", cuts0[1], c0[[isinstance(x, tuple) and x[0] == "stage" for x in c0].index(True)][2], True) in c0
    assert [x for x in c1 if isinstance(x, tuple) and x[0] == "stage"][0][3] is False
    assert c0[-2:] == ["finalize", "finalize_rows"] and c1[-2:] == ["finalize", "finalize_rows"] and "opt" not in [x[0] for x in c0 + c1 if isinstance(x, tuple)]


def test_shard_cuts_from_a_bai_index(built, tmp_path):
    """§8e: with a BAI the cut points come from the index (block offsets of its bins and linear index); they must be BGZF
    block starts, ordered, balanced about as well as the scanned cuts, and an index of another file must be refused."""
    import numpy as np
    from bamutil import write_bai
    from irfinder_b200 import synth
    from irfinder_b200.distributed import ShardPlan, _is_block_start, read_bai_block_offsets, shard_cuts, shard_cuts_from_bai
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=200, pairs=20000, readlen=100, seed=5, chunks=5, collated=False)
    bai = write_bai(meta["sorted"])
    buf = np.fromfile(meta["sorted"], dtype=np.uint8)
    offs = read_bai_block_offsets(bai)
    assert len(offs) > 20 and offs == sorted(set(offs)) and all(_is_block_start(buf, o)[0] for o in offs)
    for world in (2, 3, 8):
        cuts = shard_cuts_from_bai(bai, buf, world)
        scanned = shard_cuts(buf, world)
        assert cuts[0] == 0 and cuts[-1] == buf.size and cuts == sorted(cuts) and len(cuts) == world + 1
        assert all(c == buf.size or _is_block_start(buf, c)[0] for c in cuts[1:-1])
        gap = max(b - a for a, b in zip([0] + offs, offs + [buf.size]))           # the index is coarser than the block chain: one cut
        assert all(abs(a - b) <= gap for a, b in zip(cuts, scanned)), (cuts, scanned)   # candidate per 16 kb window that has reads
        plans = [ShardPlan.from_buffer(buf, r, world, bai=bai) for r in range(world)]
        assert sum(p.n_own for p in plans) == buf.size
    other = synth.generate(str(tmp_path / "w2"), genome="tiny", introns=200, pairs=9000, readlen=100, seed=6, chunks=3, collated=False)
    with pytest.raises(ValueError):
        shard_cuts_from_bai(write_bai(other["sorted"]), buf, 4)
