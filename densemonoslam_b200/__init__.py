"""B200-native dense RGB-D odometry: drop-in for DenseMonoSLAM / ElasticFusion `RGBDOdometry`.

The product is the C-ABI library `lib/libdmsodom.so` (hand-written CUDA for sm_100a, see
`csrc/` and `include/dmsodom/odom_c.h`); this package is the thin Python host mirror used
by the tests and the benchmark.  There is no CPU fallback: importing `odometry` without
the built library raises.
"""
__all__ = ["synth"]
