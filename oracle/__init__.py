"""CPU oracle for the batched decode hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``mlx_parallm_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may.

PARITY UNPINNED: the arithmetic of the reference lives in ``mlx==0.25.2`` /
``mlx-lm==0.24.1`` (pyproject.toml:38-39, uv.lock:527-551 of the reference),
which are not vendored under /root/reference and are not installable here (no
wheel, no network).  The reference's own tests pin no numerics
(tests/test_server_basic.py:39-52 assert status codes / JSON shapes only).  This
package therefore *restates* the published MLX op semantics (SURVEY.md
Appendix A) and the reference's model graph (mlx_parallm/models/llama.py,
qwen3.py, base.py, utils.py:315-427, sample_utils.py) in numpy; it is pinned by
hand-derived known-answer tests (tests/test_oracle_kat.py), not by outputs of
real MLX.
"""
