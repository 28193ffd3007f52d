"""CPU oracle for the UVA per-token diffusion head.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` may be imported by the product package
(``unified_video_action_b200``).  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline / ``--impl reference`` legs use it, and only as the checker
(or the timed CPU baseline), never as the thing shipped.
"""
