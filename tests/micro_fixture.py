"""A hand-derived micro fixture: 12 alignments over a 2-intron annotation where every output cell is computed by hand
(derivation in the comments).  Used to pin the oracle on CPU and, unchanged, the CUDA path on the GPU."""
import os

from bamutil import record, write_bam

REFS = [("c1", 10000), ("c0unused", 500)]

COVER = ("c1\t1000\t2000\tG1/ENSG1/clean\t0\t+\t1000\t2000\t255,0,0\t1\t990\t5\n"
         "c1\t3000\t3500\tG2/ENSG2/clean\t0\t-\t3000\t3500\t255,0,0\t2\t195,245\t5,250\n")
SJ = "c1\t1000\t2000\t+\nc1\t3000\t3500\t-\n"
RC = "c1\t1000\t+\nc1\t2000\t+\nc1\t3000\t-\nc1\t3500\t-\n"
ROI = "c1\t8000\t9000\troiA\n"


def reads():
    P, R = 0x1, 0x10   # paired, reverse
    return [
        record("A", 0, 950, "100M", 0),                      # [950,1050): intron 1 depth 1005..1049, spans point 1000 (955 < 1000 < 1046)
        record("B", 0, 900, "100M1000N50M", 0),              # junction (1000,2000) dir 1
        record("C", 0, 960, "40M1000N60M", R),               # junction (1000,2000) dir 0
        record("D", 0, 1500, "100M", 0),                     # depth 1500..1599
        record("E", 0, 1550, "100M", 0),                     # depth 1550..1649
        record("F", 0, 2950, "100M", P | 0x2 | 0x20 | 0x40, 0, 3100, 250),   # long pair (3050 < 3100): both reads kept; spans 3000
        record("F", 0, 3100, "100M", P | 0x2 | R | 0x80, 0, 2950, -250),
        record("G", 0, 3300, "100M", P | 0x2 | 0x20 | 0x40, 0, 3350, 150),   # intersect pair -> one block [3300,3450)
        record("G", 0, 3350, "100M", P | 0x2 | R | 0x80, 0, 3300, -150),
        record("H", 0, 3400, "100M", P | 0x2 | 0x20 | 0x40, 0, 3420, 100),   # short pair: read 2 inside read 1 -> [3400,3500)
        record("H", 0, 3420, "50M", P | 0x2 | R | 0x80, 0, 3400, -100),
        record("I", 0, 8100, "100M", 0),                     # inside ROI roiA
        record("J", -1, -1, "", 0x4, lseq=50),               # unmapped: skipped
        record("K", 0, 1990, "5M1000N95M", R),               # novel junction (1995,2995) dir 0; [1990,1995) + [2995,3090)
        record("L", 0, 996, "4M1000N96M", 0),                # first flank 4 < 5: junction not counted
    ]


EXPECT = {
    "IRFinder-IR-nondir.txt":
        "Chr\tStart\tEnd\tName\tNull\tStrand\tExcludedBases\tCoverage\tIntronDepth\tIntronDepth25Percentile\tIntronDepth50Percentile\tIntronDepth75Percentile\t"
        "ExonToIntronReadsLeft\tExonToIntronReadsRight\tIntronDepthFirst50bp\tIntronDepthLast50bp\tSpliceLeft\tSpliceRight\tSpliceExact\tIRratio\tWarnings\n"
        # intron 1: 990 measured positions; depth>0 on 45 (A) + 150 (D,E) + 5 (K) = 200 -> coverage 200/990; ranks [297,693) are all zero -> depth 0
        # first 50 bp [1005,1055): A covers 45 -> 0.9; last 50 bp [1945,1995): K covers 5 -> 0.1; splice left/right/exact = 2 (B,C)
        "c1\t1000\t2000\tG1/ENSG1/clean\t0\t+\t10\t0.20202\t0\t0\t0\t0\t1\t0\t0.9\t0.1\t2\t2\t2\t0\tLowCover\n"
        # intron 2: 440 measured positions; hist {0:60, 1:285, 2:95}; ranks [132,308) are all ones -> depth 1; coverage 380/440
        # first 50 bp [3005,3055): depth 2 on 45 + 1 on 5 -> 1.9; last 50 bp [3445,3495): 2 on 5 + 1 on 45 -> 1.1; no junction reads -> IRratio 1
        "c1\t3000\t3500\tG2/ENSG2/clean\t0\t-\t60\t0.863636\t1\t1\t1\t1\t1\t0\t1.9\t1.1\t0\t0\t0\t1\tLowCover\n",
    "IRFinder-JuncCount.txt":
        "JC_seqname\tstart\tend\tstrand\ttotal\tpos\tneg\n"
        "c1\t1000\t2000\t+\t2\t1\t1\n"
        "c1\t1995\t2995\t.\t1\t0\t1\n"
        "c1\t3000\t3500\t-\t0\t0\t0\n",
    "IRFinder-SpansPoint.txt":
        "SP_seqname\tcoord\ttotal\tpos\tneg\n"
        "c1\t1000\t1\t1\t0\n"      # A
        "c1\t2000\t0\t0\t0\n"
        "c1\t3000\t1\t1\t0\n"      # F read 1 (2955 < 3000 < 3046); K's block starts at 2995 so 3000 is not > 3000
        "c1\t3500\t0\t0\t0\n",
    "IRFinder-ROI.txt": "ROI_Name\ttotal\tpos\tneg\nroiA\t1\t1\t0\n",
    # fragments A,B,D,E,F,G,H,I,L forward-first (9), C and K reverse (2)
    "IRFinder-ChrCoverage.txt": "ChrCoverage_RName\ttotal\tpos\tneg\nc0unused\t0\t0\t0\nc1\t11\t9\t2\n",
}
EXPECT_STATS = {"Total reads processed": "11", "Total nucleotides": "1300", "Total singles processed": "8", "Total pairs processed": "3",
                "Short pairs": "1", "Intersect pairs": "1", "Long pairs": "1", "Skipped reads": "1", "Error reads": "0"}


def write(dirpath):
    os.makedirs(dirpath, exist_ok=True)
    p = {}
    for name, text in (("cover", COVER), ("sj", SJ), ("rc", RC), ("roi", ROI)):
        p[name] = os.path.join(dirpath, "ref-" + name + ".txt")
        with open(p[name], "w") as f:
            f.write(text)
    p["collated"] = os.path.join(dirpath, "micro.bam")
    write_bam(p["collated"], REFS, reads(), block=700)   # small BGZF blocks so records straddle them
    srt = sorted(reads(), key=lambda r: (int.from_bytes(r[4:8], "little", signed=True) & 0xFFFFFFFF, int.from_bytes(r[8:12], "little", signed=True)))
    p["sorted"] = os.path.join(dirpath, "micro.sorted.bam")
    write_bam(p["sorted"], REFS, srt, sorted_=True, block=900)
    return p
