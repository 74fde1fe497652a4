"""Frame-sharded multi-GPU execution (placeholder filled in below)."""
