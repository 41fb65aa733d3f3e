"""Build tuning variants of libob200 (different tile/stage/register settings) next to the main one."""
import sys
sys.path.insert(0, ".")
from metaopt_b200 import build as b

VARIANTS = {
    # score kernel geometry (default: 12x1 consumer warps, 2x4 fragments, BK 32, 3 stages)
    "bn64": ["SC_WM=6", "SC_WN=2", "SC_FM=4", "SC_FN=4", "SC_BK=16", "SC_STAGES=4"],
    "w42": ["SC_WM=4", "SC_WN=2", "SC_FM=8", "SC_FN=4", "SC_BK=16", "SC_STAGES=4"],
    "bk16s4": ["SC_BK=16", "SC_STAGES=4"],
    # sampler geometry (default: 32-row tiles, 4 warps, >= 8 CTAs per SM)
    "st128": ["ST_ROWS=128", "ST_THREADS=512", "ST_MIN_CTAS=2"],
    "st64": ["ST_ROWS=64", "ST_THREADS=256", "ST_MIN_CTAS=4"],
    # timing experiments only (wrong results): isolate the phases of the score kernel
    "skip2": ["SC_DEBUG_SKIP_PHASE2"],
    "skip1": ["SC_DEBUG_SKIP_PHASE1"],
    # clock64 marks around the parts of a tile (tools/prof_phases.py reads them back)
    "prof": ["SC_PROFILE_CYCLES"],
    # fit: drop one of the three kernels of a Cholesky iteration (wrong results) to see what each costs in place
    "fit_nodiag": ["FIT_DEBUG_SKIP_DIAG"],
    "fit_nopanel": ["FIT_DEBUG_SKIP_PANEL"],
    "fit_nosyrk": ["FIT_DEBUG_SKIP_SYRK"],
    # sampler with one half switched off (wrong output): which half bounds it?
    "st_db4": ["ST_NBUF=2", "ST_MIN_CTAS=4"],
    "st_db4w8": ["ST_NBUF=2", "ST_MIN_CTAS=2", "ST_THREADS=256", "ST_ROWS=64"],
    "st_rb128": ["ST_ROW_BULK_MIN=128"],
    "st7": ["ST_MIN_CTAS=7"],
    "st6": ["ST_MIN_CTAS=6"],
    "st5": ["ST_MIN_CTAS=5"],
    "st_nostore": ["ST_DEBUG_NO_STORE"],
    "st_nocompute": ["ST_DEBUG_NO_COMPUTE"],
}
if __name__ == "__main__":
    names = sys.argv[1:] or list(VARIANTS)
    for name in names:
        print(b.build(force=True, defines=VARIANTS[name], out="libob200_%s.so" % name))
