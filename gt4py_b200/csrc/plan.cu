// plan.cu -- the plan-level entry points of the C ABI (b200_plan_create / launch / destroy): see plan.inl.
#include "common.cuh"

#define B200_PLAN_SET_ERROR(msg) b200::set_error("%s", msg)
#define B200_PLAN_COLFUSE_SMEM (226 * 1024)
#include "plan.inl"
