/* TEST INFRASTRUCTURE ONLY (linked into baseline/_ref/sphexa-cuda, never into the product).
 * The reference launches a few kernels over empty ranges (e.g. buildOctreeGpu on the one-leaf tree a FocusedOctree starts
 * from: grid size 0).  CUDA answers cudaErrorInvalidConfiguration, the reference does not look, and the next thrust
 * call (CCCL 2.8 checks the last error after its own launch) throws "parallel_for failed".  With the linker's
 * --wrap=cudaLaunchKernel an empty launch becomes the no-op it was meant to be; every other launch goes through
 * untouched. */
#include <cuda_runtime.h>

extern "C" cudaError_t __real_cudaLaunchKernel(const void* func, dim3 grid, dim3 block, void** args, size_t shmem,
                                               cudaStream_t stream);

extern "C" cudaError_t __wrap_cudaLaunchKernel(const void* func, dim3 grid, dim3 block, void** args, size_t shmem,
                                               cudaStream_t stream)
{
    if (grid.x == 0 || grid.y == 0 || grid.z == 0) return cudaSuccess;
    return __real_cudaLaunchKernel(func, grid, block, args, shmem, stream);
}
