// pdl_trigger.cu -- does a CTA count as having issued griddepcontrol.launch_dependents when ANY of its threads did,
// or only when ALL of them did (or exited)?  Kernel A: every thread but thread 0 triggers at once, thread 0 spins
// `delay_us` first (mode 1), or only thread 0 triggers, late (mode 2), or all trigger at once (mode 0).  Kernel B is
// launched with the programmatic-serialization attribute and stamps its start.  nvcc -arch=sm_100a -o pdl_trigger pdl_trigger.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

__global__ void A(unsigned long long * stamps, int mode, int delay_us, int run_us) {
    const unsigned long long t0 = gt();
    if (blockIdx.x == 0 && threadIdx.x == 0) stamps[0] = t0;
    if (mode == 0 || (mode == 1 && threadIdx.x != 0)) asm volatile("griddepcontrol.launch_dependents;");
    if (threadIdx.x == 0) {
        while (gt() - t0 < (unsigned long long)delay_us * 1000) { }
        if (mode != 0) asm volatile("griddepcontrol.launch_dependents;");
    }
    while (gt() - t0 < (unsigned long long)run_us * 1000) { }
    if (blockIdx.x == 0 && threadIdx.x == 0) stamps[1] = gt();
}
__global__ void B(unsigned long long * stamps) {
    if (blockIdx.x == 0 && threadIdx.x == 0) stamps[2] = gt();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (blockIdx.x == 0 && threadIdx.x == 0) stamps[3] = gt();
}

int main() {
    unsigned long long * d; cudaMalloc(&d, 64);
    cudaStream_t s; cudaStreamCreate(&s);
    for (int mode = 0; mode < 3; mode++) {
        for (int rep = 0; rep < 3; rep++) {
            cudaMemset(d, 0, 64);
            A<<<74, 256, 0, s>>>(d, mode, 20, 60);      // half the SMs: B has room to start while A runs
            cudaLaunchConfig_t cfg = {}; cfg.gridDim = 74; cfg.blockDim = 256; cfg.stream = s;
            cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaLaunchKernelEx(&cfg, B, d);
            cudaStreamSynchronize(s);
            unsigned long long h[4]; cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
            printf("mode %d (%s): B starts %.1f us after A starts, B passes its wait at %.1f us, A ends at %.1f us\n", mode,
                   mode == 0 ? "all threads trigger at once" : mode == 1 ? "thread 0 triggers 20 us late, the others at once" : "only thread 0 triggers, 20 us late",
                   (h[2] - h[0]) / 1e3, (h[3] - h[0]) / 1e3, (h[1] - h[0]) / 1e3);
        }
    }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
