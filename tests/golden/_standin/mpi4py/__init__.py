"""Single-rank stand-in for mpi4py, used ONLY by tests/golden/make_golden.py to import the
unmodified reference in the build container (real mpi4py is not installed)."""
