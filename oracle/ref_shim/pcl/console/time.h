// Empty stand-in: odometry/src/iterative_closest_point/IterativeClosestPoint.h includes the PCL headers but declares nothing
// with their types; the ICP aligner is not on the path and is never instantiated (oracle/build_ref.py).
