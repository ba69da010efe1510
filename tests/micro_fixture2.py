"""A second hand-derived fixture: 73 single-end alignments over seven introns, built so that every rule SURVEY.md Appendix C marks as
uncertain decides at least one output cell — a 4- vs 5-base junction flank on either side, a span point exactly at blockStart+5 and at
blockEnd-4, cumulative counts exactly on the 25/50/75 % boundaries, edge windows, and one intron for each warning label.  Every
expected cell is derived by hand in the comments (default rules), and RULE_PROBES lists, per alternative rule, the cells that must
change and to what.  Used to pin the oracle on CPU and, unchanged, the CUDA path on the GPU."""
import os

from bamutil import record, write_bam

REFS = [("c1", 20000)]

# every intron has ONE measured block that leaves 5 bases out at either end (BED12: size, offset 5)
COVER = ("c1\t1000\t2000\tn1\t0\t+\t1000\t2000\t255,0,0\t1\t990\t5\n"
         "c1\t5000\t5500\tn2\t0\t+\t5000\t5500\t255,0,0\t1\t490\t5\n"
         "c1\t6000\t6110\tn3\t0\t+\t6000\t6110\t255,0,0\t1\t100\t5\n"
         "c1\t7000\t7030\tn4\t0\t+\t7000\t7030\t255,0,0\t1\t20\t5\n"
         "c1\t8000\t8100\tn5\t0\t+\t8000\t8100\t255,0,0\t1\t90\t5\n"
         "c1\t9000\t9100\tn6\t0\t+\t9000\t9100\t255,0,0\t1\t90\t5\n"
         "c1\t10000\t10100\tn7\t0\t+\t10000\t10100\t255,0,0\t1\t90\t5\n")
_INTRONS = [(1000, 2000), (5000, 5500), (6000, 6110), (7000, 7030), (8000, 8100), (9000, 9100), (10000, 10100)]
SJ = "".join(f"c1\t{s}\t{e}\t+\n" for s, e in _INTRONS)
RC = "".join(f"c1\t{p}\t+\n" for s, e in _INTRONS for p in (s, e))
ROI = "c1\t15000\t16000\troiZ\n"


def reads():
    r = []

    def add(name, pos, cig, n=1):
        for k in range(n):
            r.append(record(f"{name}_{k}", 0, pos, cig, 0))
    # n1: junction flanks.  P1 5|95 is counted, P2 4|96 and P3 96|4 are not (rule: both flanks >= 5)
    add("P1", 995, "5M1000N95M"); add("P2", 996, "4M1000N96M"); add("P3", 904, "96M1000N4M")
    # n2: span point 5000.  S1 [4995,5100): 4995+5 = 5000 is not < 5000; S2 [4900,5004): 5004-4 = 5000 is not > 5000; S3 [4994,5005): 4999 < 5000 < 5001
    add("S1", 4995, "105M"); add("S2", 4900, "104M"); add("S3", 4994, "11M")
    # n3: measured [6005,6105): depth 0 on 25 positions, 2 on 25, 4 on 25, 6 on 25 -> cumulative counts exactly on 25 / 50 / 75 %
    add("Q75", 6030, "75M", 2); add("Q50", 6055, "50M", 2); add("Q25", 6080, "25M", 2)
    # n4: depth 12 everywhere, no junction read -> LowSplicing; the reads [6995,7035) span 7030 (7030 < 7031) but not 7000 (7000 is not > 7000)
    add("L", 6995, "40M", 12)
    # n5: depth 6; 4 exact junction reads, 9 more from the same left end to 8200 -> 4 < 13 * 0.33333 -> MinorIsoform
    add("M", 8000, "100M", 6); add("Mx", 7990, "10M100N10M", 4); add("My", 7990, "10M200N10M", 9)
    # n6: depth 10, 5 exact junction reads, nothing spans the ends -> NonUniformIntronCover
    add("N", 9000, "100M", 10); add("Nx", 8990, "10M100N10M", 5)
    # n7: depth 10 from reads [9990,10110) that span both ends (10 each), 5 exact junction reads -> no warning
    add("O", 9990, "120M", 10); add("Ox", 9990, "10M100N10M", 5)
    return r


_HDR = ("Chr\tStart\tEnd\tName\tNull\tStrand\tExcludedBases\tCoverage\tIntronDepth\tIntronDepth25Percentile\tIntronDepth50Percentile\tIntronDepth75Percentile\t"
        "ExonToIntronReadsLeft\tExonToIntronReadsRight\tIntronDepthFirst50bp\tIntronDepthLast50bp\tSpliceLeft\tSpliceRight\tSpliceExact\tIRratio\tWarnings\n")
EXPECT = {
    "IRFinder-IR-nondir.txt": _HDR
        # n1: nothing reaches the measured block [1005,1995); only P1 counts for the junction; depth 0 and SpliceMax 1 -> IRratio 0
        + "c1\t1000\t2000\tn1\t0\t+\t10\t0\t0\t0\t0\t0\t0\t0\t0\t0\t1\t1\t1\t0\tLowCover\n"
        # n2: S1 covers [5005,5100) = 95 of 490 positions (0.193878); ranks [147,343) are zeros -> depth 0; first window [5005,5055) is all 1; span point 5000: S3 only
        + "c1\t5000\t5500\tn2\t0\t+\t10\t0.193878\t0\t0\t0\t0\t1\t0\t1\t0\t0\t0\t0\t0\tLowCover\n"
        # n3: 100 positions; ranks [30,70) = 20 twos + 20 fours -> 3; ties: (0+2)/2, (2+4)/2, (4+6)/2; windows [6005,6055) -> 50/50, [6055,6105) -> 250/50
        + "c1\t6000\t6110\tn3\t0\t+\t10\t0.75\t3\t1\t3\t5\t0\t0\t1\t5\t0\t0\t0\t1\tLowCover\n"
        # n4: 20 positions, all depth 12; span point 7030 twelve times
        + "c1\t7000\t7030\tn4\t0\t+\t10\t1\t12\t12\t12\t12\t0\t12\t12\t12\t0\t0\t0\t1\tLowSplicing\n"
        # n5: depth 6; left end 8000: 4 + 9 = 13, right end 8100: 4, exact 4; 6 / (6 + 13) = 0.315789
        + "c1\t8000\t8100\tn5\t0\t+\t10\t1\t6\t6\t6\t6\t0\t0\t6\t6\t13\t4\t4\t0.315789\tMinorIsoform\n"
        # n6: depth 10, 5/5/5; 10 / 15; span counts 0: 0 + 4 < 10 and 0 * 1.5 < 10
        + "c1\t9000\t9100\tn6\t0\t+\t10\t1\t10\t10\t10\t10\t0\t0\t10\t10\t5\t5\t5\t0.666667\tNonUniformIntronCover\n"
        # n7: as n6 with both span counts 10: 10 > 14 fails, 14 < 10 fails
        + "c1\t10000\t10100\tn7\t0\t+\t10\t1\t10\t10\t10\t10\t10\t10\t10\t10\t5\t5\t5\t0.666667\t-\n",
    "IRFinder-JuncCount.txt":
        "JC_seqname\tstart\tend\tstrand\ttotal\tpos\tneg\n"
        "c1\t1000\t2000\t+\t1\t1\t0\n" "c1\t5000\t5500\t+\t0\t0\t0\n" "c1\t6000\t6110\t+\t0\t0\t0\n" "c1\t7000\t7030\t+\t0\t0\t0\n"
        "c1\t8000\t8100\t+\t4\t4\t0\n" "c1\t8000\t8200\t.\t9\t9\t0\n" "c1\t9000\t9100\t+\t5\t5\t0\n" "c1\t10000\t10100\t+\t5\t5\t0\n",
    "IRFinder-SpansPoint.txt":
        "SP_seqname\tcoord\ttotal\tpos\tneg\n"
        "c1\t1000\t0\t0\t0\n" "c1\t2000\t0\t0\t0\n" "c1\t5000\t1\t1\t0\n" "c1\t5500\t0\t0\t0\n" "c1\t6000\t0\t0\t0\n" "c1\t6110\t0\t0\t0\n"
        "c1\t7000\t0\t0\t0\n" "c1\t7030\t12\t12\t0\n" "c1\t8000\t0\t0\t0\n" "c1\t8100\t0\t0\t0\n" "c1\t9000\t0\t0\t0\n" "c1\t9100\t0\t0\t0\n"
        "c1\t10000\t10\t10\t0\n" "c1\t10100\t10\t10\t0\n",
    "IRFinder-ROI.txt": "ROI_Name\ttotal\tpos\tneg\nroiZ\t0\t0\t0\n",
    "IRFinder-ChrCoverage.txt": "ChrCoverage_RName\ttotal\tpos\tneg\nc1\t73\t73\t0\n",
}
# nucleotides: 3*100 + (105+104+11) + (150+100+50) + 12*40 + (600+80+180) + (1000+100) + (1200+100) = 4560
EXPECT_STATS = {"Total reads processed": "73", "Total nucleotides": "4560", "Total singles processed": "73", "Total pairs processed": "0",
                "Skipped reads": "0", "Error reads": "0"}

# Per alternative rule: library options / oracle environment, and the cells of IRFinder-IR-nondir.txt (intron name -> {column: value}) that
# must differ from the default table, derived by hand:
#   flank 4: P2 and P3 count too -> n1 splice counts 3/3/3
#   span overhangs 4/3: S1 (4999 < 5000) and S2 (5000 < 5001) count too -> n2 ExonToIntronReadsLeft 3; the twelve n4 reads [6995,7035) now span 7000 as well (6999 < 7000)
#   no midpoint on ties: n3 percentiles 0 / 2 / 4
#   edge windows offset 3, length 40: n3 first window [6003,6043) meets the measured block in [6005,6043): 25 zeros + 13 twos -> 26/38 = 0.684211;
#       last window [6067,6107) meets it in [6067,6105): 13 fours + 25 sixes -> 202/38 = 5.31579
#   LowSplicing threshold 6: n5 (exact 4; LowSplicing is tested before MinorIsoform), n6 and n7 (exact 5) become LowSplicing
RULE_PROBES = {
    "junc_min_flank": (dict(junc_min_flank=4), {"n1": {"SpliceLeft": "3", "SpliceRight": "3", "SpliceExact": "3"}}),
    "span_overhang": (dict(span_left=4, span_right=3), {"n2": {"ExonToIntronReadsLeft": "3"}, "n4": {"ExonToIntronReadsLeft": "12"}}),
    "pct_tie": (dict(pct_tie_midpoint=0), {"n3": {"IntronDepth25Percentile": "0", "IntronDepth50Percentile": "2", "IntronDepth75Percentile": "4"}}),
    "edge_window": (dict(edge_offset=3, edge_len=40), {"n3": {"IntronDepthFirst50bp": "0.684211", "IntronDepthLast50bp": "5.31579"}}),
    "warn_lowsplice": (dict(warn_lowsplice_e6=6_000_000), {"n5": {"Warnings": "LowSplicing"}, "n6": {"Warnings": "LowSplicing"}, "n7": {"Warnings": "LowSplicing"}}),
}


def parse_ir(text):
    lines = text.rstrip("\n").split("\n")
    cols = lines[0].split("\t")
    return {f[3]: dict(zip(cols, f)) for f in (l.split("\t") for l in lines[1:])}


def write(dirpath):
    os.makedirs(dirpath, exist_ok=True)
    p = {}
    for name, text in (("cover", COVER), ("sj", SJ), ("rc", RC), ("roi", ROI)):
        p[name] = os.path.join(dirpath, "ref-" + name + ".txt")
        with open(p[name], "w") as f:
            f.write(text)
    p["collated"] = os.path.join(dirpath, "micro2.bam")
    write_bam(p["collated"], REFS, reads(), block=900)   # small BGZF blocks so records straddle them
    srt = sorted(reads(), key=lambda r: int.from_bytes(r[8:12], "little", signed=True))
    p["sorted"] = os.path.join(dirpath, "micro2.sorted.bam")
    write_bam(p["sorted"], REFS, srt, sorted_=True, block=1100)
    return p
