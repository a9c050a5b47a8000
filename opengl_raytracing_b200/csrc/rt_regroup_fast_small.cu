// rt_regroup_fast_small.cu — the regrouping megakernel, fast math, small-scene kernel, 12 warps per CTA (see rt_regroup_tu.inc).
#define RT_PARITY 0
#define RT_VARIANT_NS fast_rg_small
#define RT_RG_TU fast_small
#define RT_RG_ONLY_BIG 0
#ifndef RT_RG_WARPS_SMALL
#define RT_RG_WARPS_SMALL 12
#endif
#undef RT_PK_WARPS
#define RT_PK_WARPS RT_RG_WARPS_SMALL
#include "rt_regroup_tu.inc"
