"""Synthetic-workload tooling shared by bench.py and the tests (not part of the product)."""
