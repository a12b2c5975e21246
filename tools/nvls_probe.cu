// nvls_probe.cu -- can this box reduce in the NVSwitch (NVLS)?  Two PROCESSES, one GPU each (as under torchrun):
// every rank allocates its buffer with the virtual-memory API, rank 0 creates a multicast object and hands it to the
// other rank as a POSIX file descriptor (duplicated with pidfd_getfd: no socket needed between the ranks), every rank
// binds its buffer, and a kernel on each rank adds its contribution to ALL replicas with multimem.red.  Each rank then
// checks its own replica for the sum and times the adds.
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/nvls_probe tools/nvls_probe.cu -lcuda
//   ./nvls_probe [nranks]
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/syscall.h>
#include <sys/wait.h>
#include <unistd.h>
#include <vector>

#define CKD(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char* s_ = nullptr; cuGetErrorString(r_, &s_); fprintf(stderr, "[rank %d] %s:%d %s -> %s\n", g_rank, __FILE__, __LINE__, #x, s_ ? s_ : "?"); exit(2); } } while (0)
#define CKR(x) do { cudaError_t r_ = (x); if (r_ != cudaSuccess) { fprintf(stderr, "[rank %d] %s:%d %s -> %s\n", g_rank, __FILE__, __LINE__, #x, cudaGetErrorString(r_)); exit(2); } } while (0)
static int g_rank = 0;

__global__ void k_fill(int* p, size_t n, int v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void k_mc_add(int* mc, size_t n, int v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        asm volatile("multimem.red.relaxed.sys.global.add.s32 [%0], %1;" ::"l"(mc + i), "r"(v) : "memory");
}
// reduce-scatter + all-gather through the switch: this rank pulls the sum over all replicas of its slice of `src`
// (multimem.ld_reduce: the switch adds) and multicasts it into every replica of `dst` (multimem.st)
__global__ void k_mc_rs_ag(const unsigned long long* mc_src, unsigned long long* mc_dst, size_t n64, int rank, int nranks) {
    const size_t per = (n64 + nranks - 1) / nranks, lo = per * rank, hi = lo + per < n64 ? lo + per : n64;
    for (size_t i = lo + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
        unsigned long long v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(mc_src + i) : "memory");
        asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(mc_dst + i), "l"(v) : "memory");
    }
}
__global__ void k_check(const int* p, size_t n, int expect, unsigned long long* bad) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        if (p[i] != expect) atomicAdd(bad, 1ull);
}

struct Pipe { int rd, wr; };
static void put(int fd, const void* p, size_t n) { if (write(fd, p, n) != (ssize_t)n) { perror("write"); exit(3); } }
static void get(int fd, void* p, size_t n) { size_t g = 0; while (g < n) { ssize_t r = read(fd, (char*)p + g, n - g); if (r <= 0) { perror("read"); exit(3); } g += r; } }

int main(int argc, char** argv) {
    const int nranks = argc > 1 ? atoi(argv[1]) : 2;
    // pipes: to_child[r] (rank 0 -> rank r), to_root[r] (rank r -> rank 0)
    std::vector<Pipe> down(nranks), up(nranks);
    for (int r = 1; r < nranks; ++r) { int a[2], b[2]; if (pipe(a) || pipe(b)) return 3; down[r] = {a[0], a[1]}; up[r] = {b[0], b[1]}; }
    for (int r = 1; r < nranks; ++r) { pid_t p = fork(); if (p == 0) { g_rank = r; break; } }
    auto barrier = [&]() {
        char c = 1;
        if (g_rank == 0) { for (int r = 1; r < nranks; ++r) get(up[r].rd, &c, 1); for (int r = 1; r < nranks; ++r) put(down[r].wr, &c, 1); }
        else { put(up[g_rank].wr, &c, 1); get(down[g_rank].rd, &c, 1); }
    };
    CKD(cuInit(0));
    int ndev = 0; CKD(cuDeviceGetCount(&ndev));
    if (ndev < nranks) { if (g_rank == 0) printf("only %d device(s) visible for %d ranks\n", ndev, nranks); return 0; }
    CKR(cudaSetDevice(g_rank));
    CKR(cudaFree(0));
    CUdevice dev; CKD(cuDeviceGet(&dev, g_rank));
    int mc_ok = 0; CKD(cuDeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
    if (g_rank == 0) printf("multicast supported: %d\n", mc_ok);
    if (!mc_ok) return 0;

    const size_t bytes = 512ull << 20;      // one 512^3 int32 grid
    CUmulticastObjectProp mp; memset(&mp, 0, sizeof mp);
    mp.numDevices = nranks; mp.size = bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR; mp.flags = 0;
    size_t gmin = 0, grec = 0;
    CKD(cuMulticastGetGranularity(&gmin, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    CKD(cuMulticastGetGranularity(&grec, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (g_rank == 0) printf("multicast granularity: minimum %zu, recommended %zu\n", gmin, grec);

    // the rank's own physical memory
    CUmemAllocationProp ap; memset(&ap, 0, sizeof ap);
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED; ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ap.location.id = g_rank;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t agran = 0; CKD(cuMemGetAllocationGranularity(&agran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (g_rank == 0) printf("allocation granularity %zu\n", agran);
    CUmemGenericAllocationHandle mem; CKD(cuMemCreate(&mem, bytes, &ap, 0));
    CUdeviceptr uc = 0; CKD(cuMemAddressReserve(&uc, bytes, grec, 0, 0)); CKD(cuMemMap(uc, bytes, 0, mem, 0));
    CUmemAccessDesc ad; ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ad.location.id = g_rank; ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CKD(cuMemSetAccess(uc, bytes, &ad, 1));

    // the multicast object: created by rank 0, handed over as a file descriptor
    CUmemGenericAllocationHandle mc;
    if (g_rank == 0) {
        CKD(cuMulticastCreate(&mc, &mp));
        int fd = -1; CKD(cuMemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
        int msg[2] = {(int)getpid(), fd};
        for (int r = 1; r < nranks; ++r) put(down[r].wr, msg, sizeof msg);
    } else {
        int msg[2]; get(down[g_rank].rd, msg, sizeof msg);
        int pidfd = (int)syscall(434 /* pidfd_open */, msg[0], 0);
        if (pidfd < 0) { perror("pidfd_open"); return 4; }
        int fd = (int)syscall(438 /* pidfd_getfd */, pidfd, msg[1], 0);
        if (fd < 0) { perror("pidfd_getfd"); return 4; }
        CKD(cuMemImportFromShareableHandle(&mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
        close(fd); close(pidfd);
    }
    CKD(cuMulticastAddDevice(mc, dev));
    barrier();                                  // every device added before anybody binds
    CKD(cuMulticastBindMem(mc, 0, mem, 0, bytes, 0));
    CUdeviceptr mv = 0; CKD(cuMemAddressReserve(&mv, bytes, grec, 0, 0)); CKD(cuMemMap(mv, bytes, 0, mc, 0));
    CKD(cuMemSetAccess(mv, bytes, &ad, 1));
    barrier();
    if (g_rank == 0) printf("multicast object bound and mapped on %d ranks\n", nranks);

    const size_t n = bytes / 4;
    k_fill<<<592, 1024>>>((int*)uc, n, 0); CKR(cudaDeviceSynchronize());
    barrier();
    // every rank adds (rank + 1) to the first `box` words through the switch
    const size_t box = 238328;                  // 62^3 cells
    cudaEvent_t e0, e1; CKR(cudaEventCreate(&e0)); CKR(cudaEventCreate(&e1));
    k_mc_add<<<148, 1024>>>((int*)mv, box, g_rank + 1); CKR(cudaDeviceSynchronize());
    barrier();
    unsigned long long* bad; CKR(cudaMallocManaged(&bad, 8)); *bad = 0;
    k_check<<<148, 1024>>>((const int*)uc, box, nranks * (nranks + 1) / 2, bad); CKR(cudaDeviceSynchronize());
    printf("[rank %d] words of the 62^3 box that are not the sum over ranks: %llu\n", g_rank, *bad);
    barrier();
    const int reps = 200;
    CKR(cudaEventRecord(e0));
    for (int i = 0; i < reps; ++i) k_mc_add<<<148, 1024>>>((int*)mv, box, 1);
    CKR(cudaEventRecord(e1)); CKR(cudaDeviceSynchronize());
    float ms = 0; CKR(cudaEventElapsedTime(&ms, e0, e1));
    printf("[rank %d] multimem.red of a 62^3 box (%.2f MB): %.2f us per launch, all ranks adding at once\n", g_rank, box * 4 / 1e6, ms * 1e3 / reps);
    barrier();
    *bad = 0;
    k_check<<<148, 1024>>>((const int*)uc, box, nranks * (nranks + 1) / 2 + nranks * reps, bad); CKR(cudaDeviceSynchronize());
    printf("[rank %d] after the timed adds: %llu wrong words\n", g_rank, *bad);
    barrier();
    // reduce-scatter + all-gather: src = words [0, box) (every replica holds `fill` there), dst = words [box2, box2 + box)
    {
        const size_t box2 = 64ull << 20;          // 256 MiB further on
        k_fill<<<592, 1024>>>((int*)uc, box, g_rank + 1); k_fill<<<592, 1024>>>((int*)uc + box2, box, 0); CKR(cudaDeviceSynchronize());
        barrier();
        k_mc_rs_ag<<<148, 1024>>>((const unsigned long long*)mv, (unsigned long long*)((int*)mv + box2), box / 2, g_rank, nranks); CKR(cudaDeviceSynchronize());
        barrier();
        *bad = 0;
        k_check<<<148, 1024>>>((const int*)uc + box2, box, nranks * (nranks + 1) / 2, bad); CKR(cudaDeviceSynchronize());
        printf("[rank %d] reduce-scatter + all-gather (ld_reduce + multimem.st): %llu wrong words\n", g_rank, *bad);
        barrier();
        CKR(cudaEventRecord(e0));
        for (int i = 0; i < reps; ++i) k_mc_rs_ag<<<148, 1024>>>((const unsigned long long*)mv, (unsigned long long*)((int*)mv + box2), box / 2, g_rank, nranks);
        CKR(cudaEventRecord(e1)); CKR(cudaDeviceSynchronize());
        CKR(cudaEventElapsedTime(&ms, e0, e1));
        printf("[rank %d] ld_reduce of a 1/%d slice of the 62^3 box + multimem.st of the result: %.2f us per launch, all ranks at once\n", g_rank, nranks, ms * 1e3 / reps);
        barrier();
        // for scale: an empty kernel launched the same way
        CKR(cudaEventRecord(e0));
        for (int i = 0; i < reps; ++i) k_fill<<<148, 1024>>>((int*)uc, 0, 0);
        CKR(cudaEventRecord(e1)); CKR(cudaDeviceSynchronize());
        CKR(cudaEventElapsedTime(&ms, e0, e1));
        printf("[rank %d] empty kernel: %.2f us per launch\n", g_rank, ms * 1e3 / reps);
        barrier();
    }
    if (g_rank == 0) for (int r = 1; r < nranks; ++r) wait(nullptr);
    return 0;
}
