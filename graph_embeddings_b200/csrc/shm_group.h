// Host-side plumbing for the ranks of ONE node: a POSIX shared-memory segment with a barrier and one
// mailbox per rank.  The block-rotation trainer moves its data through CUDA IPC peer memory (dist.cu);
// what is left for a communicator are a few hundred bytes per epoch (IPC handles, the scalar loss, x_max,
// nnz).  ncclCommInitRank + the first collective cost 3.2 s at 8 ranks (2.1 s + 1.0 s measured, more than
// the 0.76 s of the 8 epochs of the C4 bench), so on a single node these bytes go through this segment
// and NCCL is only brought up when a plane that needs it is selected (GE_DIST_ROT=nccl|pipe, or IPC is
// unavailable).  No reference counterpart (the reference is one JVM: Optimizer.java:59-63 slices threads).
#ifndef GE_SHM_GROUP_H
#define GE_SHM_GROUP_H
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

struct ShmGroup {
    static constexpr uint32_t MAGIC = 0x47455348u;  // "GESH"
    static constexpr size_t SLOT = 4096;            // mailbox bytes per rank
    struct Hdr {
        std::atomic<uint32_t> magic, count, gen, attached;
        uint32_t world, pad[11];
    };
    Hdr* h = nullptr;
    unsigned char* slots = nullptr;
    size_t bytes = 0;
    int rank = 0, world = 1;
    double timeout_s = 120.0;
    std::string name;

    // rank 0 creates the segment, the others wait for it; the name comes from the job's 128-byte unique id
    bool open(const void* id128, int rank_, int world_, std::string& err) {
        rank = rank_; world = world_;
        uint64_t hsh = 1469598103934665603ull;
        for (int k = 0; k < 128; k++) hsh = (hsh ^ ((const unsigned char*)id128)[k]) * 1099511628211ull;
        char nm[64];
        snprintf(nm, sizeof nm, "/graphemb_%016llx", (unsigned long long)hsh);
        name = nm;
        bytes = sizeof(Hdr) + SLOT * (size_t)world;
        const char* to = getenv("GE_SHM_TIMEOUT");
        if (to) timeout_s = atof(to);
        int fd = -1;
        const auto t0 = std::chrono::steady_clock::now();
        if (rank == 0) {
            shm_unlink(nm);  // a stale segment of a crashed job with the same id cannot exist, but be safe
            fd = shm_open(nm, O_CREAT | O_EXCL | O_RDWR, 0600);
            if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) { err = std::string("shm_open/ftruncate ") + nm + " failed"; if (fd >= 0) ::close(fd); return false; }
        } else {
            for (;;) {
                fd = shm_open(nm, O_RDWR, 0600);
                struct stat sb;
                if (fd >= 0 && fstat(fd, &sb) == 0 && (size_t)sb.st_size >= bytes) break;
                if (fd >= 0) ::close(fd);
                if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) {
                    err = std::string("shared-memory segment ") + nm + " of rank 0 did not appear (ranks on different nodes?)";
                    return false;
                }
                usleep(200);
            }
        }
        void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        ::close(fd);
        if (p == MAP_FAILED) { err = "mmap of the shared-memory segment failed"; return false; }
        h = reinterpret_cast<Hdr*>(p);
        slots = reinterpret_cast<unsigned char*>(p) + sizeof(Hdr);
        if (rank == 0) {
            h->count.store(0); h->gen.store(0); h->attached.store(0); h->world = (uint32_t)world;
            h->magic.store(MAGIC, std::memory_order_release);
        } else {
            while (h->magic.load(std::memory_order_acquire) != MAGIC) {
                if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) { err = "shared-memory segment never initialised"; return false; }
                usleep(100);
            }
            if (h->world != (uint32_t)world) { err = "shared-memory segment belongs to a job of another size"; return false; }
        }
        h->attached.fetch_add(1);
        if (!barrier(err)) return false;
        if (rank == 0) shm_unlink(nm);  // everybody has it mapped: the name can go, nothing leaks if a rank dies
        return true;
    }
    void close() {
        if (h) munmap(h, bytes);
        h = nullptr; slots = nullptr;
    }
    // sense-reversing barrier; false on timeout (a peer died or never came)
    bool barrier(std::string& err) {
        const uint32_t g = h->gen.load(std::memory_order_acquire);
        if (h->count.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)world) {
            h->count.store(0, std::memory_order_relaxed);
            h->gen.store(g + 1, std::memory_order_release);
            return true;
        }
        const auto t0 = std::chrono::steady_clock::now();
        for (unsigned spin = 0;; spin++) {
            if (h->gen.load(std::memory_order_acquire) != g) return true;
            if (spin < 4096) continue;
            if ((spin & 1023) == 0 && std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) {
                err = "rank " + std::to_string(rank) + ": a peer did not reach the barrier (GE_SHM_TIMEOUT)";
                return false;
            }
            usleep(20);
        }
    }
    // every rank contributes n <= SLOT bytes; all[k*n .. ) = rank k's
    bool allgather(const void* mine, void* all, size_t n, std::string& err) {
        if (n > SLOT) { err = "shm allgather: message too large"; return false; }
        memcpy(slots + SLOT * (size_t)rank, mine, n);
        if (!barrier(err)) return false;
        for (int k = 0; k < world; k++) memcpy((unsigned char*)all + n * (size_t)k, slots + SLOT * (size_t)k, n);
        return barrier(err);  // nobody overwrites a mailbox somebody still reads
    }
};
#endif
