// Stand-in for the reference's lukas_kanade/include/lukas_kanade/lukas_kanade.h without ForwardAdditive.h (not on the
// alignment path).  The headers named here are the reference's own (-I <reference>/lukas_kanade/src).
#ifndef REF_SHIM_LUKAS_KANADE_H
#define REF_SHIM_LUKAS_KANADE_H
#include "InverseCompositional.h"
#include "InverseCompositionalStacked.h"
#include "Warp.h"
#endif
