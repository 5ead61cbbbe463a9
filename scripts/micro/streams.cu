// micro test: does work on one stream proceed while another stream waits (kernel spin / cuStreamWaitValue32)?
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>
__global__ void spin(volatile unsigned* f, unsigned v) { unsigned n = 0; while (*f != v) { __nanosleep(100); if (++n > (1u << 26)) { printf("spin timeout\n"); return; } } }
__global__ void setf(volatile unsigned* f, unsigned v) { *f = v; __threadfence_system(); }
typedef CUresult (*WaitFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned);
int main(int argc, char** argv)
{
    int mode = argc > 1 ? atoi(argv[1]) : 0;
    unsigned* f; cudaMalloc(&f, 64); cudaMemset(f, 0, 64);
    cudaStream_t a, b; cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking);
    cudaDeviceSynchronize();
    if (mode == 0) spin<<<1, 32, 0, a>>>(f, 1);
    else {
        void* p = 0; cudaDriverEntryPointQueryResult q;
        cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q);
        printf("entry %p q %d\n", p, (int)q);
        CUresult r = ((WaitFn)p)((CUstream)a, (CUdeviceptr)f, 1, CU_STREAM_WAIT_VALUE_EQ);
        printf("wait enqueue rc %d\n", (int)r);
        setf<<<1, 32, 0, a>>>(f + 8, 7);
    }
    if (mode == 2) usleep(500000);
    setf<<<1, 32, 0, b>>>(f, 1);
    printf("enqueued\n"); fflush(stdout);
    cudaError_t e = cudaStreamSynchronize(b);
    printf("b done: %s\n", cudaGetErrorString(e)); fflush(stdout);
    e = cudaStreamSynchronize(a);
    printf("a done: %s\n", cudaGetErrorString(e));
    unsigned h[16]; cudaMemcpy(h, f, 64, cudaMemcpyDeviceToHost); printf("flags %u %u\n", h[0], h[8]);
    return 0;
}
