// Host logic of the multi-rank set-up in picb200_peer_connect (csrc/pic_nvls.h), exercised without a GPU:
// the barrier in a POSIX shared-memory segment between PROCESSES (fork, as under torchrun), a rank that reports a
// failure (the others must give up at once, not after the time-out), the time-out itself, and a file descriptor
// duplicated out of another process with pidfd_getfd.  Exit code 0 and "ok" on success.
#include "../../particle-simulation_b200/csrc/pic_nvls.h"

#include <cstdio>
#include <cstring>
#include <sys/wait.h>

using namespace picb200;

static int run_ranks(const char* name, int nranks, int failing_rank, int missing_rank, int timeout_s, int expect_ok) {
    shm_unlink(name);
    ShmBarrier* root = shm_barrier_open(name, true);
    if (!root) { perror("shm create"); return 1; }
    for (int r = 0; r < nranks; ++r) {
        if (r == missing_rank) continue;
        if (fork() == 0) {
            ShmBarrier* b = shm_barrier_open(name, false);
            if (!b) _exit(9);
            if (r == failing_rank) { shm_barrier_fail(b); _exit(7); }
            const auto t0 = std::chrono::steady_clock::now();
            const bool ok0 = shm_barrier_wait(b, 0, nranks, timeout_s);
            const bool ok1 = ok0 && shm_barrier_wait(b, 1, nranks, timeout_s);
            const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (failing_rank >= 0 && s > 1.0) _exit(8);          // a reported failure must not cost the time-out
            _exit(ok1 ? 0 : 1);
        }
    }
    int good = 0, bad = 0, other = 0, st = 0;
    while (wait(&st) > 0) {
        const int c = WIFEXITED(st) ? WEXITSTATUS(st) : 99;
        if (c == 0) ++good; else if (c == 1 || c == 7) ++bad; else ++other;
    }
    munmap(root, 4096);
    shm_unlink(name);
    const int present = nranks - (missing_rank >= 0 ? 1 : 0);
    if (other) { fprintf(stderr, "%s: %d ranks ended unexpectedly\n", name, other); return 1; }
    if (expect_ok ? good != present : bad != present) { fprintf(stderr, "%s: good %d bad %d of %d\n", name, good, bad, present); return 1; }
    return 0;
}

int main() {
    char name[64];
    snprintf(name, sizeof name, "/picb200-test-%d", (int)getpid());
    if (run_ranks(name, 8, -1, -1, 20, 1)) return 1;             // eight ranks meet twice
    if (run_ranks(name, 4, 2, -1, 20, 0)) return 1;              // rank 2 reports a failure: everybody gives up within a second
    if (run_ranks(name, 3, -1, 1, 1, 0)) return 1;               // rank 1 never calls: the others time out (1 s here, 60 s in the library)
    // a descriptor out of another process: the child opens a pipe and writes through the stolen write end
    int p[2];
    if (pipe(p)) return 1;
    const pid_t parent = getpid();
    if (fork() == 0) {
        const int fd = steal_fd((long long)parent, p[1]);
        if (fd < 0) _exit(5);
        _exit(write(fd, "x", 1) == 1 ? 0 : 6);
    }
    int st = 0;
    wait(&st);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
        if (WIFEXITED(st) && WEXITSTATUS(st) == 5) { printf("ok (pidfd_getfd not permitted here)\n"); return 0; }
        fprintf(stderr, "steal_fd child failed (%d)\n", WIFEXITED(st) ? WEXITSTATUS(st) : -1);
        return 1;
    }
    char c = 0;
    if (read(p[0], &c, 1) != 1 || c != 'x') { fprintf(stderr, "nothing came through the stolen descriptor\n"); return 1; }
    printf("ok\n");
    return 0;
}
