// Stand-in for the reference's core/include/core/core.h: the same headers FROM THE REFERENCE TREE (found through -I
// <reference>/core/src), minus Point3D.h and Trajectory.h, which the alignment path does not use.  Test infrastructure.
#ifndef REF_SHIM_CORE_H
#define REF_SHIM_CORE_H
#include "Camera.h"
#include "Exceptions.h"
#include "Feature2D.h"
#include "Frame.h"
#include "Kernel2d.h"
#include "PoseWithCovariance.h"
#include "algorithm.h"
#include "image_transform.h"
#include "macros.h"
#include "types.h"
#endif
