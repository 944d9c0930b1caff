#!/usr/bin/env python
"""Recipe for ``baseline/_ref``: a git-ignored, UNMODIFIED copy of the reference's hot-path files.

    python baseline/make_ref.py            (build container only: needs /root/reference)

The reference (WohimLee/NICE-SLAM-annotation) is plain Python without setup.py / pyproject.toml, so there is
nothing to ``pip install --target baseline/_ref``; the files are copied byte for byte instead (``src/`` without
the offline ``tools/`` and ``configs/``).  ``baseline/_ref`` is listed in .gitignore (no reference source enters
the history) but not in .gpurunignore, so it travels to the GPU box with the snapshot, where

* ``bench.py`` times the reference's own ``Renderer.render_batch_ray`` / ``Tracker.optimize_cam_in_batch`` on the
  B200 (``torch_cuda_baseline``, ``kind: "reference"``) and on the host cores (``--impl reference``,
  ``cpu_baseline``), and
* ``tests/test_reference_seam.py`` runs the reference's own ``optimize_cam_in_batch`` / ``optimize_map`` source on
  top of this package's Renderer / decoders (the drop-in seam of INTEGRATION.md section 1).

``__graft_entry__.build()`` calls ``make()`` whenever /root/reference exists.  Nothing under baseline/_ref is
imported by the product package.
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC_DEFAULT = os.environ.get("NICE_SLAM_REFERENCE", "/root/reference")
# what the hot path imports plus the modules those files import at module level
FILES = ["src/__init__.py", "src/common.py", "src/config.py", "src/Mapper.py", "src/Tracker.py", "src/NICE_SLAM.py",
         "src/conv_onet/__init__.py", "src/conv_onet/config.py", "src/conv_onet/models/__init__.py",
         "src/conv_onet/models/decoder.py", "src/utils/Renderer.py", "src/utils/Visualizer.py",
         "src/utils/datasets.py", "src/utils/Logger.py", "src/utils/Mesher.py", "LICENSE"]
TREES = ["configs"]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def make(src=SRC_DEFAULT, dst=DST, quiet=False):
    """Copy the files; returns the number of files written (0 when everything was already identical)."""
    if not os.path.isdir(os.path.join(src, "src")):
        raise FileNotFoundError(f"{src} is not a checkout of the reference")
    todo = list(FILES)
    for t in TREES:
        for root, _, names in os.walk(os.path.join(src, t)):
            todo += [os.path.relpath(os.path.join(root, n), src) for n in names]
    written, lines = 0, []
    for rel in sorted(todo):
        a, b = os.path.join(src, rel), os.path.join(dst, rel)
        if not os.path.exists(a):
            continue
        sha = _sha(a)
        lines.append(f"{sha}  {rel}")
        if os.path.exists(b) and _sha(b) == sha:
            continue
        os.makedirs(os.path.dirname(b), exist_ok=True)
        shutil.copyfile(a, b)
        written += 1
    with open(os.path.join(dst, "MANIFEST.sha256"), "w") as f:
        f.write("\n".join(lines) + "\n")
    if not quiet:
        print(f"baseline/_ref: {len(lines)} files ({written} written) from {src}")
    return written


def available(dst=DST):
    return os.path.exists(os.path.join(dst, "src", "utils", "Renderer.py"))


if __name__ == "__main__":
    make(*(sys.argv[1:2] or [SRC_DEFAULT]))
