"""Data parallelism: flat parameter/gradient buckets, the fused P2P all-reduce(+optimizer) engine, launch helpers."""
