// pic_nvls.h -- host side of the NVSwitch reduction (NVLS) used by the multi-GPU exchange (pic_peer.cuh).
//
// With a multicast object bound to the summed charge grids of all ranks, ONE `multimem.red` adds a word into every
// rank's grid: the reduction fans out inside the switch, a rank sends its box once instead of once per peer (7 MB ->
// 1 MB of NVLink egress per rank and Poisson step at 8 GPUs; the reference has no multi-GPU path, src/pic.cu:394).
// The grids then have to come from the virtual-memory API (cuMemCreate), and the multicast object has to reach every
// rank: rank 0 creates it and exports a POSIX file descriptor, the others duplicate that descriptor out of rank 0's
// process with pidfd_getfd (one process per GPU under torchrun; no socket between the ranks is needed) or, inside one
// process, use the handle as it is.  "Every device added before any memory is bound" needs a barrier between the
// ranks inside picb200_peer_connect: a counter in a POSIX shared-memory segment named in rank 0's handle.
//
// The driver API is reached through cudaGetDriverEntryPoint, so libpicb200.so keeps loading on machines without
// libcuda (the CPU-side ABI tests).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <string>
#include <thread>

namespace picb200 {

struct DrvApi {
#define PIC_DRV(name) decltype(&name) name##_ = nullptr;
    PIC_DRV(cuDeviceGet)
    PIC_DRV(cuDeviceGetAttribute)
    PIC_DRV(cuGetErrorString)
    PIC_DRV(cuMemCreate)
    PIC_DRV(cuMemRelease)
    PIC_DRV(cuMemAddressReserve)
    PIC_DRV(cuMemAddressFree)
    PIC_DRV(cuMemMap)
    PIC_DRV(cuMemUnmap)
    PIC_DRV(cuMemSetAccess)
    PIC_DRV(cuMemExportToShareableHandle)
    PIC_DRV(cuMemImportFromShareableHandle)
    PIC_DRV(cuMulticastCreate)
    PIC_DRV(cuMulticastAddDevice)
    PIC_DRV(cuMulticastBindMem)
    PIC_DRV(cuMulticastGetGranularity)
#undef PIC_DRV
    bool ok = false;
};

inline DrvApi* drv_api() {
    static DrvApi api;
    static bool tried = false;
    if (tried) return api.ok ? &api : nullptr;
    tried = true;
    bool all = true;
    auto get = [&](const char* name, void** fn) {
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !*fn) { all = false; cudaGetLastError(); }
    };
#define PIC_DRV(name) get(#name, reinterpret_cast<void**>(&api.name##_));
    PIC_DRV(cuDeviceGet)
    PIC_DRV(cuDeviceGetAttribute)
    PIC_DRV(cuGetErrorString)
    PIC_DRV(cuMemCreate)
    PIC_DRV(cuMemRelease)
    PIC_DRV(cuMemAddressReserve)
    PIC_DRV(cuMemAddressFree)
    PIC_DRV(cuMemMap)
    PIC_DRV(cuMemUnmap)
    PIC_DRV(cuMemSetAccess)
    PIC_DRV(cuMemExportToShareableHandle)
    PIC_DRV(cuMemImportFromShareableHandle)
    PIC_DRV(cuMulticastCreate)
    PIC_DRV(cuMulticastAddDevice)
    PIC_DRV(cuMulticastBindMem)
    PIC_DRV(cuMulticastGetGranularity)
#undef PIC_DRV
    api.ok = all;
    return api.ok ? &api : nullptr;
}

inline std::string drv_error(DrvApi* d, CUresult r) {
    const char* s = nullptr;
    if (d && d->cuGetErrorString_) d->cuGetErrorString_(r, &s);
    return s ? std::string(s) : ("CUresult " + std::to_string((int)r));
}

// barrier between the ranks of one node during set-up: two counters in a POSIX shared-memory segment
struct ShmBarrier {
    std::atomic<int> count[4];
};
inline ShmBarrier* shm_barrier_open(const char* name, bool create) {
    int fd = shm_open(name, create ? (O_CREAT | O_EXCL | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) return nullptr;
    if (create && ftruncate(fd, 4096) != 0) { close(fd); shm_unlink(name); return nullptr; }
    void* p = mmap(nullptr, 4096, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    return p == MAP_FAILED ? nullptr : static_cast<ShmBarrier*>(p);      // (a fresh segment is all zero)
}
// count[3]: ranks whose set-up failed -- nobody waits for them (they say so and everybody gives up at once)
inline void shm_barrier_fail(ShmBarrier* b) { b->count[3].fetch_add(1, std::memory_order_acq_rel); }
inline bool shm_barrier_wait(ShmBarrier* b, int which, int nranks, int timeout_s = 60) {
    b->count[which].fetch_add(1, std::memory_order_acq_rel);
    const auto t0 = std::chrono::steady_clock::now();
    while (b->count[which].load(std::memory_order_acquire) < nranks) {
        if (b->count[3].load(std::memory_order_acquire) > 0) return false;
        if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(timeout_s)) return false;
        std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
    return b->count[3].load(std::memory_order_acquire) == 0;
}

// a file descriptor of another process of this user, duplicated into this one (Linux >= 5.6)
inline int steal_fd(long long pid, int fd) {
    const int pidfd = (int)syscall(434 /* pidfd_open */, (pid_t)pid, 0);
    if (pidfd < 0) return -1;
    const int mine = (int)syscall(438 /* pidfd_getfd */, pidfd, fd, 0);
    close(pidfd);
    return mine;
}

}  // namespace picb200
