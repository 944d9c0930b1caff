"""CPU oracle for the NICE-SLAM render hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``nice_slam_annotation_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` use it, and only as the checker or the
reported CPU baseline, never as the thing shipped.

Parity status: the reference (WohimLee/NICE-SLAM-annotation) ships no tests or golden
vectors for this path, so the oracle is pinned against OUTPUTS OF THE
REFERENCE ITSELF, produced in the build container by ``tests/golden/make_golden.py``
(which imports ``/root/reference`` unmodified) and committed under ``tests/golden/``.
``tests/test_oracle_golden.py`` checks the oracle against those vectors.
"""
from .nice_oracle import *  # noqa: F401,F403
