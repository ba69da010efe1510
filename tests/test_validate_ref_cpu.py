"""CPU suite: the BuildRef-format validator on generated and deliberately damaged reference sets."""
import os
import shutil

from conftest import ROOT  # noqa: F401  (puts the repo on sys.path)


def _refs(tmp_path):
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=120, pairs=500, readlen=100, seed=3, chunks=2, collated=False)
    return os.path.dirname(meta["cover"])


def _damaged(src, tmp_path, name, edit):
    d = str(tmp_path / ("bad_" + name.replace(".", "_")))
    shutil.copytree(src, d)
    p = os.path.join(d, name)
    lines = open(p).read().split("\n")
    open(p, "w").write("\n".join(edit(lines)))
    return d


def test_generated_reference_sets_are_consistent(built, tmp_path):
    from irfinder_b200.validate_ref import main, validate_refdir
    d = _refs(tmp_path)
    assert validate_refdir(d) == []
    assert main([d]) == 0


def test_damaged_reference_sets_are_reported(built, tmp_path):
    from irfinder_b200.validate_ref import main, validate_refdir
    d = _refs(tmp_path)

    def has(problems, sev, text):
        return any(s == sev and text in m for s, m in problems)

    def blocks_overlap(ls):
        i = next(k for k, l in enumerate(ls) if l and int(l.split("\t")[9]) >= 2)     # swap the first two blocks of a multi-block record
        f = ls[i].split("\t")
        a, b = f[10].split(","), f[11].split(",")
        a[0], a[1], b[0], b[1] = a[1], a[0], b[1], b[0]
        f[10], f[11] = ",".join(a), ",".join(b)
        return ls[:i] + ["\t".join(f)] + ls[i + 1:]

    cases = [
        ("ref-cover.bed", lambda ls: [ls[0].rsplit("\t", 1)[0]] + ls[1:], "error", "12 BED columns"),
        ("ref-cover.bed", lambda ls: ["\t".join(ls[0].split("\t")[:9] + ["9"] + ls[0].split("\t")[10:])] + ls[1:], "error", "blockCount 9"),
        ("ref-cover.bed", blocks_overlap, "error", "starts before block"),
        ("ref-cover.bed", lambda ls: [ls[1], ls[0]] + ls[2:], "warning", "not sorted"),
        ("ref-cover.bed", lambda ls: [ls[0]] + ls, "warning", "listed twice"),
        ("ref-sj.ref", lambda ls: ls[1:], "error", "has no junction line"),
        ("ref-sj.ref", lambda ls: ["\t".join(ls[0].split("\t")[:3] + ["?"])] + ls[1:], "error", "strand '?'"),
        ("ref-read-continues.ref", lambda ls: ls[1:], "error", "is missing from ref-read-continues.ref"),
        ("ref-ROI.bed", lambda ls: ["\t".join([ls[0].split("\t")[0], "500", "100", "backwards"])] + ls, "error", "before start"),
    ]
    for name, edit, sev, text in cases:
        bad = _damaged(d, tmp_path, name, edit)
        problems = validate_refdir(bad)
        assert has(problems, sev, text), (name, text, problems[:5])
        assert main([bad]) == (1 if sev == "error" else 0)
        shutil.rmtree(bad)
