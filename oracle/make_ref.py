"""Recipe for ``oracle/_ref/``: the reference's own hot-path files, copied UNMODIFIED from ``/root/reference`` so that the timed CPU
baseline of ``bench.py`` (``cpu_baseline``, ``--impl reference``) can run the reference's code itself (``kind: "reference"``) on
the GPU box, where ``/root/reference`` does not exist.  ``oracle/_ref/`` is git-ignored (never part of the history) but travels
with the ``gpurun`` snapshot, like the built ``.so``.  The reference is pure Python: there is nothing to compile.

    python oracle/make_ref.py        # no-op when /root/reference is absent
"""
import os
import shutil

REF = "/root/reference"
FILES = ("rmvd/models/blocks/planesweep_corr.py", "rmvd/models/blocks/utils.py")
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def make():
    if not os.path.isdir(os.path.join(REF, "rmvd")):
        return False
    for rel in FILES:
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF, rel), dst)
    return True


if __name__ == "__main__":
    print("oracle/_ref written" if make() else "/root/reference absent: nothing done")
