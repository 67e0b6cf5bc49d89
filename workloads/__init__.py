"""Synthetic inputs (matrices, vectors, element meshes) for the BASELINE.json configurations.  Pure numpy: shared by
bench.py (both arms), tests/ and tools/ — importing it loads no native library."""
