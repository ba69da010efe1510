"""Validator for the BuildRef-format inputs of the counting core: ref-cover.bed, ref-sj.ref,
ref-read-continues.ref and ref-ROI.bed of one <RefDir>/IRFinder directory.

The library's loaders (csrc/refs.cpp) reject lines they cannot parse; this tool adds the checks that need the whole file
or several files at once: BED12 block consistency, sort order, duplicates, and agreement between the three files (every
measured intron is an annotated junction, both its splice sites are span points).  Host-only, not on the counting path.

    python -m irfinder_b200.validate_ref <RefDir>/IRFinder        # exit status 1 when an error was found
"""
import os
import sys


def _lines(path):
    with open(path, "r") as f:
        for n, line in enumerate(f, 1):
            line = line.rstrip("\n").rstrip("\r")
            if line:
                yield n, line.split("\t")


def _int(f, what, where, out):
    if not f.isdigit():
        out.append(("error", f"{where}: {what} '{f}' is not a non-negative integer"))
        return None
    return int(f)


def check_cover(path, out):
    """BED12 rows; returns {(chr, start, end): strand}."""
    introns, prev = {}, None
    name = os.path.basename(path)
    for n, f in _lines(path):
        w = f"{name} line {n}"
        if len(f) < 12:
            out.append(("error", f"{w}: expected 12 BED columns, found {len(f)}"))
            continue
        s, e, bc = _int(f[1], "start", w, out), _int(f[2], "end", w, out), _int(f[9], "blockCount", w, out)
        if s is None or e is None or bc is None:
            continue
        if e <= s:
            out.append(("error", f"{w}: end {e} is not after start {s}"))
        if f[5] not in "+-" or len(f[5]) != 1:
            out.append(("error", f"{w}: strand '{f[5]}' is neither + nor -"))
        sizes, starts = [x for x in f[10].split(",") if x], [x for x in f[11].split(",") if x]
        if len(sizes) != bc or len(starts) != bc:
            out.append(("error", f"{w}: blockCount {bc} but {len(sizes)} sizes and {len(starts)} starts"))
            continue
        pos = s
        for i, (a, b) in enumerate(zip(sizes, starts)):
            a, b = _int(a, "block size", w, out), _int(b, "block start", w, out)
            if a is None or b is None:
                break
            if s + b < pos and i:
                out.append(("error", f"{w}: block {i + 1} starts before block {i} ends"))
            if s + b + a > e:
                out.append(("error", f"{w}: block {i + 1} reaches beyond the record's end"))
            if a == 0:
                out.append(("warning", f"{w}: block {i + 1} is empty"))
            pos = s + b + a
        key = (f[0], s, e)
        if key in introns:
            out.append(("warning", f"{w}: intron {f[0]}:{s}-{e} is listed twice"))
        introns[key] = f[5]
        if prev is not None and prev[0] == f[0] and (s, e) < prev[1:]:
            out.append(("warning", f"{w}: not sorted by start within {f[0]} (BuildRef writes this file sorted)"))
        prev = (f[0], s, e)
    return introns


def check_sj(path, out):
    junc = {}
    name = os.path.basename(path)
    for n, f in _lines(path):
        w = f"{name} line {n}"
        if len(f) < 4:
            out.append(("error", f"{w}: expected chr, start, end, strand"))
            continue
        s, e = _int(f[1], "start", w, out), _int(f[2], "end", w, out)
        if s is None or e is None:
            continue
        if e <= s:
            out.append(("error", f"{w}: end {e} is not after start {s}"))
        if f[3] not in ("+", "-", "."):
            out.append(("error", f"{w}: strand '{f[3]}' is not one of + - ."))
        key = (f[0], s, e)
        if key in junc and f[3] in junc[key]:
            out.append(("warning", f"{w}: junction {f[0]}:{s}-{e} {f[3]} is listed twice"))
        junc.setdefault(key, set()).add(f[3])
    return junc


def check_points(path, out):
    pts = set()
    name = os.path.basename(path)
    for n, f in _lines(path):
        w = f"{name} line {n}"
        if len(f) < 2:
            out.append(("error", f"{w}: expected chr, position[, strand]"))
            continue
        p = _int(f[1], "position", w, out)
        if p is not None:
            pts.add((f[0], p))
    return pts


def check_roi(path, out):
    name = os.path.basename(path)
    seen = set()
    for n, f in _lines(path):
        w = f"{name} line {n}"
        if len(f) < 4:
            out.append(("error", f"{w}: expected chr, start, end, name"))
            continue
        s, e = _int(f[1], "start", w, out), _int(f[2], "end", w, out)
        if s is not None and e is not None and e < s:
            out.append(("error", f"{w}: end {e} before start {s}"))
        if f[3] in seen:
            out.append(("warning", f"{w}: ROI name '{f[3]}' is used twice (rows are reported per line, names only label them)"))
        seen.add(f[3])


def validate(cover, sj, read_continues, roi=None):
    """Returns [(severity, message)], severity in {"error", "warning"}; an empty list means the set is consistent."""
    out = []
    introns, junc, pts = check_cover(cover, out), check_sj(sj, out), check_points(read_continues, out)
    if roi and os.path.basename(roi) != "NULL" and os.path.exists(roi):
        check_roi(roi, out)
    cname, sname, pname = os.path.basename(cover), os.path.basename(sj), os.path.basename(read_continues)
    for (c, s, e), strand in introns.items():
        st = junc.get((c, s, e))
        if st is None:
            out.append(("error", f"{cname}: intron {c}:{s}-{e} has no junction line in {sname} (its SpliceExact column could never count)"))
        elif strand not in st and "." not in st:
            out.append(("warning", f"{cname}: intron {c}:{s}-{e} is on {strand} but {sname} lists the junction on {','.join(sorted(st))}"))
        for p in (s, e):
            if (c, p) not in pts:
                out.append(("error", f"{cname}: splice site {c}:{p} of intron {c}:{s}-{e} is missing from {pname} (ExonToIntronReads would stay 0)"))
    return out


def validate_refdir(refdir):
    j = lambda n: os.path.join(refdir, n)
    return validate(j("ref-cover.bed"), j("ref-sj.ref"), j("ref-read-continues.ref"), j("ref-ROI.bed"))


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) != 1:
        print(__doc__.strip().split("\n\n")[-1], file=sys.stderr)
        return 2
    problems = validate_refdir(argv[0])
    for sev, msg in problems:
        print(f"{sev}: {msg}")
    errors = sum(1 for sev, _ in problems if sev == "error")
    print(f"{errors} error(s), {len(problems) - errors} warning(s)")
    return 1 if errors else 0


if __name__ == "__main__":
    sys.exit(main())
