"""Shared helpers of the drop-in step scripts (same argv and files as the reference's programs)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_ENGINE = None
# set by run_pipeline.py when it runs the steps inside its own process: keep BEDs in HBM between steps and
# write the re-projected BED files in the background (run_pipeline.py flushes them before it exits)
KEEP_RESIDENT = False


def engine():
    """One Engine per process (device from SALSA_B200_DEVICE, default 0)."""
    global _ENGINE
    if _ENGINE is None:
        from salsa_b200.engine import Engine
        _ENGINE = Engine(int(os.environ.get('SALSA_B200_DEVICE', '0')))
    return _ENGINE


def clean_pass():
    """The CUDA BED rewrite for salsa_b200.host.clean.clean."""
    from salsa_b200 import reproject
    return reproject.clean_pass(engine())


def re_pass():
    """The CUDA per-scaffold RE proration for salsa_b200.host.reproject.update_bed."""
    from salsa_b200 import reproject
    return reproject.re_pass(engine())


def bed_pass():
    """The CUDA BED pass for salsa_b200.host.reproject.update_bed (no CPU alternative here)."""
    from salsa_b200 import reproject
    return reproject.bed_pass(engine(), background=KEEP_RESIDENT)


_LINKS = {}


def remember_links(path, node_names, u, v, w, links):
    """make_links.py -> fast_scaled_scores.py inside one process: the link list as arrays, valid while the file
    it was written to is unchanged."""
    st = os.stat(path)
    _LINKS.clear()
    _LINKS[os.path.realpath(path)] = (st.st_size, st.st_mtime_ns, (node_names, u, v, w, links))


def cached_links(path):
    hit = _LINKS.get(os.path.realpath(path))
    if hit is None:
        return None
    st = os.stat(path)
    return hit[2] if (hit[0], hit[1]) == (st.st_size, st.st_mtime_ns) else None
