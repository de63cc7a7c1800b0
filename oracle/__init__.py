"""TEST INFRASTRUCTURE ONLY -- CPU oracles for the B200 hot path.

Nothing under ``oracle/`` is on the product path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and only as the checker (or the timed CPU baseline), never as the
thing shipped.  ``neural_magic_vllm_b200`` never imports this package.
"""
