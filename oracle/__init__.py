"""TEST INFRASTRUCTURE — CPU oracle for the Whisper hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package. The product (``whisper_tflite_b200``)
never does, and fails loudly when its CUDA library is missing.
"""
