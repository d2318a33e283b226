"""Data-parallel machinery: synchronize, DistributedOptimizer, the device world (symmetric
heap + NVLink kernels) and the fused engine. DP is the only strategy the reference has
(SURVEY §2.2), so it is the only one here."""
