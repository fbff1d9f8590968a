"""BASELINES, not product: restated step structure of the reference's own GPU
backend, timed beside ours by ``bench.py --impl refcuda-port``."""
