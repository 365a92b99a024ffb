// Stand-in for <THC/THCAtomics.cuh>: sm_60+ has native atomicAdd(double*, double).
#pragma once
