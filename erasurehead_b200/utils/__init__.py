"""Host utilities: reference-compatible file I/O, result writers, checkpoints, clocks."""
