// ref_reduce_wrap.cpp -- TEST INFRASTRUCTURE.  Gives the checker a linkable handle on the reference's OWN
// reduceCPU<T> (Kahan gold, Samples/2_Concepts_and_Techniques/reduction/reduction.cpp:167-180) and
// getNumBlocksAndThreads (:204-240).  Both live inside the sample's driver translation unit, so this file
// includes that TU from where it lies under /root/reference (oracle/build.py passes its directory with -I;
// nothing is copied) with its main() renamed, and exports plain-C wrappers.  The sample's GPU entry point
// reduce<T>() (reduction.h:32, defined in reduction_kernel.cu) is satisfied by an aborting stub: the
// checker never launches the reference's kernels from here.
#include <cstdlib>

#define main reduction_sample_main_unused
#include "reduction.cpp"
#undef main

template <class T> void reduce(int, int, int, int, T *, T *) { abort(); }
template void reduce<int>(int, int, int, int, int *, int *);
template void reduce<float>(int, int, int, int, float *, float *);
template void reduce<double>(int, int, int, int, double *, double *);

extern "C" int    ref_reduceCPU_int(int *data, int size) { return reduceCPU<int>(data, size); }
extern "C" float  ref_reduceCPU_float(float *data, int size) { return reduceCPU<float>(data, size); }
extern "C" double ref_reduceCPU_double(double *data, int size) { return reduceCPU<double>(data, size); }
// needs a CUDA device (it reads the device properties): GPU tests only
extern "C" void ref_getNumBlocksAndThreads(int whichKernel, int n, int maxBlocks, int maxThreads, int *blocks,
                                           int *threads)
{
    getNumBlocksAndThreads(whichKernel, n, maxBlocks, maxThreads, *blocks, *threads);
}
