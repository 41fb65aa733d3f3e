"""Build tuning variants of libob200 (different tile/stage/register settings) next to the main one."""
import sys
sys.path.insert(0, ".")
from metaopt_b200 import build as b

VARIANTS = {
    "bn64": ["SC_WM=6", "SC_WN=2", "SC_FM=4", "SC_FN=4", "SC_BK=16", "SC_STAGES=4"],
    "bn32s4": ["SC_STAGES=4"],
    "st128": ["ST_ROWS=128", "ST_THREADS=512", "ST_MIN_CTAS=2"],
    "st64": ["ST_ROWS=64", "ST_THREADS=256", "ST_MIN_CTAS=4"],
    "st32w8": ["ST_ROWS=32", "ST_THREADS=256", "ST_MIN_CTAS=4"],
    "bk32s2": ["SC_BK=32", "SC_STAGES=2"],
    "bk16s5": ["SC_BK=16", "SC_STAGES=5"],
    "bk8s8": ["SC_BK=8", "SC_STAGES=8"],
    "r224": ["SC_CONSUMER_REGS=224"],
    "p1s3": ["P1_STAGES=3"],
    "w42": ["SC_WM=4", "SC_WN=2", "SC_FM=8", "SC_FN=4"],
    "w62bk32": ["SC_BK=32", "SC_STAGES=2"],
    "w62bk32s3": ["SC_BK=32", "SC_STAGES=3"],
    "skip2": ["SC_DEBUG_SKIP_PHASE2"],
    "skip1": ["SC_DEBUG_SKIP_PHASE1"],
}
if __name__ == "__main__":
    names = sys.argv[1:] or list(VARIANTS)
    for name in names:
        print(b.build(force=True, defines=VARIANTS[name], out="libob200_%s.so" % name))
