"""CPU oracle for the spectral hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker / the CPU arm being timed.
The product package (``task-aware-audio-coding-perceptual_b200/tac_spectral``) never imports it
and raises when its CUDA library is missing.
"""
