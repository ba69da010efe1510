"""ctypes binding of include/irfinder_b200.h and the IRFinderCore convenience class."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    return os.path.join(_HERE, "libirfinder_b200.so")


class IRFError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"irfinder_b200 error {code}: {msg}")
        self.code = code


TABLES = dict(stats=0, spans=1, roi=2, chr=3, junctions=4, depth=5, rows_nondir=6, rows_dir=7, fragments=8, timing=9, diff=10)

ROW_DTYPE = np.dtype([("excluded", "<i8"), ("positions", "<u8"), ("coverage", "<f8"), ("depth", "<f8"), ("p25", "<f8"), ("p50", "<f8"),
                      ("p75", "<f8"), ("first50", "<f8"), ("last50", "<f8"), ("irratio", "<f8"), ("sp_left", "<u4"), ("sp_right", "<u4"),
                      ("jc_left", "<u4"), ("jc_right", "<u4"), ("jc_exact", "<u4"), ("warning", "<i4")])
JUNC_DTYPE = np.dtype([("refid", "<i4"), ("start", "<u4"), ("end", "<u4"), ("neg", "<u4"), ("pos", "<u4")])
_TABLE_DTYPES = {0: np.dtype("<u8"), 1: np.dtype("<u4"), 2: np.dtype("<u4"), 3: np.dtype("<u4"), 4: JUNC_DTYPE, 5: np.dtype("<i4"),
                 6: ROW_DTYPE, 7: ROW_DTYPE, 8: np.dtype("<u4"), 9: np.dtype("<f8")}
STAT_NAMES = ["singles", "pairs", "short", "intersect", "long", "skipped", "error", "nucleotides", "records"]

# every symbol include/irfinder_b200.h declares: name -> (restype, argtypes)
_vp, _sz, _i, _cp = C.c_void_p, C.c_size_t, C.c_int, C.c_char_p
SYMBOLS = {
    "irf_create": (_i, [C.POINTER(_vp), C.POINTER(_i), _i]),
    "irf_destroy": (None, [_vp]),
    "irf_last_error": (_cp, [_vp]),
    "irf_set_option": (_i, [_vp, _cp, _cp]),
    "irf_load_ref": (_i, [_vp, _cp, _cp, _cp, _cp]),
    "irf_run_bam": (_i, [_vp, _cp, _cp]),
    "irf_run_buffer": (_i, [_vp, _vp, _sz]),
    "irf_run_stream": (_i, [_vp, _i, _i]),
    "irf_write_outputs": (_i, [_vp, _cp]),
    "irf_stdout_text": (_cp, [_vp]),
    "irf_directionality": (_i, [_vp]),
    "irf_get_table": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_sz)]),
    "irf_reset": (_i, [_vp]),
    "irf_read_header": (_i, [_vp, _vp, _sz]),
    "irf_stage": (_i, [_vp, _vp, _sz, _sz, _i]),
    "irf_count": (_i, [_vp]),
    "irf_frame_info": (_i, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "irf_finalize": (_i, [_vp]),
    "irf_get_device_table": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_sz), C.POINTER(_i)]),
    "irf_export_junctions": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_sz)]),
    "irf_import_junctions": (_i, [_vp, _i, _vp, _vp, _sz]),
    "irf_export_halo": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz), C.POINTER(_vp), C.POINTER(_sz)]),
    "irf_import_halo": (_i, [_vp, _vp, _sz, _vp, _sz, _i, _i]),
    "irf_first_record": (_i, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "irf_owner_plan": (_i, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int64), _i, _i]),
    "irf_export_foreign_depth": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_sz)]),
    "irf_import_depth": (_i, [_vp, _vp, _vp, _sz]),
    "irf_finalize_rows": (_i, [_vp]),
    "irf_depth_scan_static": (_i, [_vp, _vp, _sz, _vp, C.c_char_p, _sz]),
}
_lib = None


def load_library():
    """Loads libirfinder_b200.so; raises ImportError (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise ImportError(f"{p} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "or `make -C irfinder_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def _b(s):
    return None if s is None else os.fsencode(s)


class IRFinderCore:
    """Host mirror of the reference's `irfinder` main: loadRef x4 -> processAll -> WriteOutput x6."""

    def __init__(self, device=0, **options):
        """`device`: one GPU ordinal, or a list of them (an ordinal may repeat): irf_run_bam / irf_run_buffer then cut the BAM into one
        byte range per entry, count them side by side inside this process and merge into the first context (csrc/multi.cu)."""
        self.lib = load_library()
        self.h = _vp()
        ids = [int(d) for d in device] if isinstance(device, (list, tuple)) else [int(device)]
        self.devices = ids
        dev = (_i * len(ids))(*ids)
        rc = self.lib.irf_create(C.byref(self.h), dev, len(ids))
        if rc != 0:
            raise IRFError(rc, "irf_create failed: no usable CUDA device (this library has no CPU path)")
        for k, v in options.items():
            self.set_option(k, v)

    def close(self):
        if getattr(self, "h", None):
            self.lib.irf_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise IRFError(rc, (self.lib.irf_last_error(self.h) or b"").decode(errors="replace"))

    def set_option(self, key, value):
        self._ck(self.lib.irf_set_option(self.h, key.encode(), str(int(value)).encode()))

    def load_ref(self, cover_bed, sj_ref, read_continues_ref, roi_bed=None):
        self._ck(self.lib.irf_load_ref(self.h, _b(cover_bed), _b(sj_ref), _b(read_continues_ref), _b(roi_bed) if roi_bed else b"NULL"))

    def load_ref_dir(self, ref_dir):
        """<RefDir>/IRFinder/ref-*.{bed,ref} layout used by `IRFinder -m BAM -r <RefDir>`."""
        d = os.path.join(ref_dir, "IRFinder")
        roi = os.path.join(d, "ref-ROI.bed")
        self.load_ref(os.path.join(d, "ref-cover.bed"), os.path.join(d, "ref-sj.ref"), os.path.join(d, "ref-read-continues.ref"), roi if os.path.exists(roi) else None)

    def run_bam(self, path):
        self._ck(self.lib.irf_run_bam(self.h, _b(path), None))

    def run_buffer(self, ptr, nbytes):
        self._ck(self.lib.irf_run_buffer(self.h, _vp(ptr), nbytes))

    def run_bytes(self, data):
        buf = np.frombuffer(data, dtype=np.uint8)
        self.run_buffer(buf.ctypes.data, buf.size)

    def run_stream(self, fd, inflated):
        self._ck(self.lib.irf_run_stream(self.h, fd, int(bool(inflated))))

    def write_outputs(self, outdir):
        os.makedirs(outdir, exist_ok=True)
        self._ck(self.lib.irf_write_outputs(self.h, _b(outdir)))

    @property
    def stdout_text(self):
        return self.lib.irf_stdout_text(self.h).decode()

    @property
    def directionality(self):
        return self.lib.irf_directionality(self.h)

    def table(self, which):
        code = TABLES[which] if isinstance(which, str) else which
        p, n = _vp(), _sz()
        self._ck(self.lib.irf_get_table(self.h, code, C.byref(p), C.byref(n)))
        dt = _TABLE_DTYPES[code]
        if n.value == 0:
            return np.zeros(0, dtype=dt)
        raw = (C.c_uint8 * (n.value * dt.itemsize)).from_address(p.value)
        return np.frombuffer(raw, dtype=dt).copy()

    def stats(self):
        return dict(zip(STAT_NAMES, (int(x) for x in self.table("stats"))))

    def timing(self):
        t = self.table("timing")
        keys = ["inflate_wall_ms", "inflate_kernel_ms", "framing_ms", "counting_ms", "finalize_ms", "count_total_ms", "host_scan_ms", "output_ms", "launches", "inflate_launches", "sm_inflate_ms", "sm_inflate_share", "tok_launches_total", "adjacent_pairs"]
        return dict(zip(keys, (float(x) for x in t)))

    # ---- staged / sharded operation
    def reset(self):
        self._ck(self.lib.irf_reset(self.h))

    def read_header(self, ptr, nbytes):
        self._ck(self.lib.irf_read_header(self.h, _vp(ptr), nbytes))

    def stage(self, ptr, nbytes_own, nbytes_total=None, has_header=True):
        self._ck(self.lib.irf_stage(self.h, _vp(ptr), nbytes_own, nbytes_own if nbytes_total is None else nbytes_total, int(has_header)))

    def count(self):
        self._ck(self.lib.irf_count(self.h))

    def frame_info(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.irf_frame_info(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def finalize(self):
        self._ck(self.lib.irf_finalize(self.h))

    def device_table(self, which):
        code = TABLES[which] if isinstance(which, str) else which
        p, n, eb = _vp(), _sz(), _i()
        self._ck(self.lib.irf_get_device_table(self.h, code, C.byref(p), C.byref(n), C.byref(eb)))
        return p.value or 0, n.value, eb.value

    def export_junctions(self, which):
        k, c, n = _vp(), _vp(), _sz()
        self._ck(self.lib.irf_export_junctions(self.h, which, C.byref(k), C.byref(c), C.byref(n)))
        return k.value or 0, c.value or 0, n.value

    def import_junctions(self, which, keys_ptr, counts_ptr, n):
        self._ck(self.lib.irf_import_junctions(self.h, which, _vp(keys_ptr), _vp(counts_ptr), n))

    def export_halo(self):
        b, nb, o, nr = _vp(), _sz(), _vp(), _sz()
        self._ck(self.lib.irf_export_halo(self.h, C.byref(b), C.byref(nb), C.byref(o), C.byref(nr)))
        return b.value or 0, nb.value, o.value or 0, nr.value

    # ---- region-owned depth (sharded runs)
    def first_record(self):
        a, b = C.c_int32(), C.c_int32()
        self._ck(self.lib.irf_first_record(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def owner_plan(self, refids, positions, rank):
        n = len(refids)
        ra = (C.c_int32 * n)(*[int(x) for x in refids])
        pa = (C.c_int64 * n)(*[int(x) for x in positions])
        self._ck(self.lib.irf_owner_plan(self.h, ra, pa, n, rank))

    def export_foreign_depth(self):
        i, v, n = _vp(), _vp(), _sz()
        self._ck(self.lib.irf_export_foreign_depth(self.h, C.byref(i), C.byref(v), C.byref(n)))
        return i.value or 0, v.value or 0, n.value

    def import_depth(self, idx_ptr, val_ptr, n):
        self._ck(self.lib.irf_import_depth(self.h, _vp(idx_ptr), _vp(val_ptr), n))

    def finalize_rows(self):
        self._ck(self.lib.irf_finalize_rows(self.h))

    def import_halo(self, bytes_ptr, nbytes, offsets_ptr, nrec, owner, n_owners):
        self._ck(self.lib.irf_import_halo(self.h, _vp(bytes_ptr), nbytes, _vp(offsets_ptr), nrec, owner, n_owners))


def depth_scan_static(diff_ptr, depth_ptr, n, stream=None):
    """Allocation-free depth scan (irf_depth_scan_static): for hosts whose GPU memory is fully pre-partitioned, or that
    cannot afford the device-wide stall of cudaMalloc/cudaFree inside their loop.  Device pointers in, nothing returned;
    the work is queued on `stream`.  At most 2**27 counters per call; calls on one GPU must not overlap (shared scratch)."""
    lib = load_library()
    err = C.create_string_buffer(512)
    rc = lib.irf_depth_scan_static(_vp(diff_ptr), _vp(depth_ptr), int(n), _vp(stream) if stream else None, err, len(err))
    if rc != 0:
        raise IRFError(rc, err.value.decode(errors="replace"))
