"""slam_b200 -- B200-native (sm_100a) batched two-class 2-D ICP: the ccicp2d hot
path of servos/SLAM behind a C ABI (include/ccicp2d_b200.h).  See DESIGN.md."""
__all__ = ["capi", "scangen", "build"]
