"""Ships the UNMODIFIED reference to the GPU box -- TEST / BASELINE INFRASTRUCTURE ONLY.

``/root/reference`` exists in the build container but not on the GPU box.  This recipe copies the two
pure-Python packages the hot path's callers live in,

    /root/reference/dreamer        ->  oracle/_ref/dreamer
    /root/reference/rlpyt/rlpyt    ->  oracle/_ref/rlpyt/rlpyt

into ``oracle/_ref/`` (git-ignored: no reference source ever enters the history; NOT gpurun-ignored, so
the copy travels with the tree exactly like the built ``.so``).  ``oracle/ref_harness.py`` falls back to
it when ``/root/reference`` is absent.  It is used for three things only:

  * ``bench.py --impl reference`` / ``cpu_baseline``: the reference's own modules timed on the host cores,
  * ``bench.py``'s ``reference_cuda`` block: the reference's eager CUDA path on the same B200 (the
    north star's ">= 10x" denominator),
  * ``tests/test_reference_algo_gpu.py``: the reference's own ``Dreamer.loss`` run on the drop-in modules.

The product (``dreamer_pytorch_b200/``) never imports it.

    python -m oracle.fetch_ref            # run by __graft_entry__.build() when /root/reference exists
"""
from __future__ import annotations

import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SRC = os.environ.get("DREAMER_REFERENCE_SRC", "/root/reference")

_PAIRS = (("dreamer", "dreamer"), (os.path.join("rlpyt", "rlpyt"), os.path.join("rlpyt", "rlpyt")))


def fetch(src: str = SRC, dest: str = DEST) -> bool:
    """Returns True if ``dest`` holds a copy afterwards (False: no source checkout and no earlier copy)."""
    if not os.path.isdir(os.path.join(src, "dreamer")):
        return os.path.isdir(os.path.join(dest, "dreamer"))
    ignore = shutil.ignore_patterns("__pycache__", "*.pyc", "*.so", "*.c", "*.pyx")
    for a, b in _PAIRS:
        out = os.path.join(dest, b)
        if os.path.isdir(out):
            shutil.rmtree(out)
        shutil.copytree(os.path.join(src, a), out, ignore=ignore)
    with open(os.path.join(dest, "README"), "w") as f:
        f.write("Unmodified copy of juliusfrost/dreamer-pytorch (dreamer/, rlpyt/rlpyt/) made by oracle/fetch_ref.py.\n"
                "Git-ignored; test and baseline infrastructure only.\n")
    return True


if __name__ == "__main__":
    ok = fetch()
    print(f"oracle/_ref: {'ready' if ok else 'no reference checkout found'}")
    sys.exit(0 if ok else 1)
