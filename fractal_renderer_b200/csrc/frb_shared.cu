// frb_shared.cu — frb_shared_frame_*: one host image shared by the processes of a one-process-per-GPU job.
//
// The reference collects the rows of one image into one Vec<RGB> (src/lib.rs:256-267).  With one process per GPU that "collect"
// would be a gather; here it is no step at all: the image lives in shared memory (a file in /dev/shm), every rank maps it and registers the
// mapping with CUDA (page-locked, portable), and each rank's frb_render_bands writes its interleaved bands straight into it —
// by the streaming finisher over PCIe while the iterations run.  The header page in front of the image carries a
// sense-reversing barrier so that "the whole frame is on the host" costs a few cache-line transfers, not a collective.
#include "../../include/frb.h"

#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/statvfs.h>
#include <unistd.h>

int frb_internal_fail(int code, const char *fmt, ...);  // frb_api.cu: sets frb_last_error() of the calling thread

namespace {

constexpr uint64_t kMagic = 0x46524253484d3031ull;  // "FRBSHM01"
constexpr size_t kHeaderBytes = 4096;

struct Header {
    std::atomic<uint64_t> magic;   // written last by rank 0: the frame is ready to be attached
    uint64_t bytes;
    uint32_t ranks;
    uint32_t pad;
    alignas(64) std::atomic<uint32_t> arrived;
    alignas(64) std::atomic<uint32_t> sense;
    alignas(64) std::atomic<uint32_t> attached;  // ranks that have mapped the frame (rank 0 may unlink the name once all have)
};
static_assert(sizeof(Header) <= kHeaderBytes, "header page");
static_assert(std::atomic<uint32_t>::is_always_lock_free && std::atomic<uint64_t>::is_always_lock_free, "atomics in shared memory");

struct Frame {
    std::string name;
    Header *hdr = nullptr;
    uint8_t *data = nullptr;
    size_t map_bytes = 0;
    uint32_t ranks = 1, rank = 0;
    uint32_t local_sense = 0;
    bool pinned = false;
    int fd = -1;  // ranks != 0 keep the file open: fstat tells whether the frame they mapped is still the one the name refers to
};

// True when the file behind `fd` has been unlinked, or `path` now names another file: a rank that attached to the left-over frame
// of a crashed job just before this job's rank 0 removed it would otherwise render into memory nobody reads.
bool is_stale(int fd, const std::string &path)
{
    struct stat mine, named;
    if (fd < 0 || fstat(fd, &mine) != 0) return false;
    if (mine.st_nlink == 0) return true;
    return stat(path.c_str(), &named) == 0 && (named.st_ino != mine.st_ino || named.st_dev != mine.st_dev);
}

void cpu_relax()
{
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
}

}  // namespace

extern "C" {

FrbSharedFrame *frb_shared_frame_open(const char *name, size_t bytes, uint32_t ranks, uint32_t rank)
{
    if (!name || name[0] != '/' || strlen(name) < 2 || strchr(name + 1, '/')) {
        frb_internal_fail(FRB_ERR_INVALID, "shared frame name must look like \"/frb-job\"");
        return nullptr;
    }
    if (ranks == 0 || rank >= ranks) {
        frb_internal_fail(FRB_ERR_INVALID, "rank %u of %u", rank, ranks);
        return nullptr;
    }
    const size_t page = 4096;
    const size_t map_bytes = kHeaderBytes + ((bytes + page - 1) / page) * page;
    // The frame is a file in a memory-backed directory: /dev/shm where it has room (containers often cap it at 64 MB), else /tmp.
    static const char *const kDirs[] = {"/dev/shm", "/tmp"};
    int fd = -1;
    std::string path;
    if (rank == 0) {
        for (const char *d : kDirs) unlink((std::string(d) + name).c_str());  // a stale frame of a crashed job
        for (const char *d : kDirs) {
            struct statvfs vfs;
            if (statvfs(d, &vfs) != 0 || (uint64_t)vfs.f_bavail * vfs.f_frsize < (uint64_t)map_bytes + (64u << 20)) continue;
            path = std::string(d) + name;
            fd = open(path.c_str(), O_CREAT | O_EXCL | O_RDWR | O_CLOEXEC, 0600);
            if (fd >= 0) break;
        }
        if (fd < 0) {
            frb_internal_fail(FRB_ERR_INVALID, "no room for a %zu-byte shared frame %s in /dev/shm or /tmp: %s", map_bytes, name, strerror(errno));
            return nullptr;
        }
        if (posix_fallocate(fd, 0, (off_t)map_bytes) != 0 && ftruncate(fd, (off_t)map_bytes) != 0) {
            frb_internal_fail(FRB_ERR_INVALID, "cannot size %s to %zu bytes: %s", path.c_str(), map_bytes, strerror(errno));
            close(fd);
            unlink(path.c_str());
            return nullptr;
        }
    } else {
        // rank 0 creates; wait for the name, then for its size, then (below) for the magic
        const auto t0 = std::chrono::steady_clock::now();
        while (fd < 0) {
            for (const char *d : kDirs) {
                path = std::string(d) + name;
                fd = open(path.c_str(), O_RDWR | O_CLOEXEC);
                if (fd < 0) continue;
                struct stat st;
                if (fstat(fd, &st) == 0 && (size_t)st.st_size >= map_bytes) break;
                close(fd);
                fd = -1;
            }
            if (fd >= 0) break;
            if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(30)) {
                frb_internal_fail(FRB_ERR_INVALID, "shared frame %s did not appear within 30 s (is rank 0 running?)", name);
                return nullptr;
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(2));
        }
    }
    void *m = mmap(nullptr, map_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (rank == 0) {
        close(fd);
        fd = -1;
    }
    if (m == MAP_FAILED) {
        frb_internal_fail(FRB_ERR_INVALID, "mmap(%s, %zu): %s", name, map_bytes, strerror(errno));
        if (rank == 0) unlink(path.c_str());
        if (fd >= 0) close(fd);
        return nullptr;
    }
    Frame *f = new (std::nothrow) Frame();
    if (!f) {
        munmap(m, map_bytes);
        if (fd >= 0) close(fd);
        return nullptr;
    }
    f->fd = fd;
    f->name = path;
    f->hdr = static_cast<Header *>(m);
    f->data = static_cast<uint8_t *>(m) + kHeaderBytes;
    f->map_bytes = map_bytes;
    f->ranks = ranks;
    f->rank = rank;
    if (rank == 0) {
        // fresh shared memory is zero-filled: arrived = sense = attached = 0
        f->hdr->bytes = bytes;
        f->hdr->ranks = ranks;
        f->hdr->magic.store(kMagic, std::memory_order_release);
    } else {
        const auto t0 = std::chrono::steady_clock::now();
        while (f->hdr->magic.load(std::memory_order_acquire) != kMagic) {
            if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(30)) {
                frb_internal_fail(FRB_ERR_INVALID, "shared frame %s was never initialised by rank 0", name);
                munmap(m, map_bytes);
                close(f->fd);
                delete f;
                return nullptr;
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(1));
        }
        // The magic may be that of a crashed job's frame under the same name, opened before this job's rank 0 got to remove it:
        // if the name has moved on, attach again (to the new file).  The barrier re-checks (a removal that comes later still).
        static thread_local int reattach = 0;
        if (is_stale(f->fd, path) && reattach < 100) {
            munmap(m, map_bytes);
            close(f->fd);
            delete f;
            std::this_thread::sleep_for(std::chrono::milliseconds(10));
            ++reattach;
            FrbSharedFrame *again = frb_shared_frame_open(name, bytes, ranks, rank);
            --reattach;
            return again;
        }
        if (f->hdr->bytes != bytes || f->hdr->ranks != ranks) {
            frb_internal_fail(FRB_ERR_INVALID, "shared frame %s is %llu bytes for %u ranks, not %zu for %u", name,
                              (unsigned long long)f->hdr->bytes, f->hdr->ranks, bytes, ranks);
            munmap(m, map_bytes);
            close(f->fd);
            delete f;
            return nullptr;
        }
    }
    // touch every page now (first-touch faults belong to set-up, not to the first frame), then pin if a device is there
    for (size_t o = 0; o < map_bytes - kHeaderBytes; o += page) reinterpret_cast<volatile uint8_t *>(f->data)[o] |= 0;
    if (frb_device_count() > 0 && bytes > 0) {
        if (cudaHostRegister(f->data, map_bytes - kHeaderBytes, cudaHostRegisterPortable) == cudaSuccess) f->pinned = true;
        else cudaGetLastError();
    }
    f->hdr->attached.fetch_add(1, std::memory_order_acq_rel);
    return reinterpret_cast<FrbSharedFrame *>(f);
}

uint8_t *frb_shared_frame_data(FrbSharedFrame *h)
{
    Frame *f = reinterpret_cast<Frame *>(h);
    return f ? f->data : nullptr;
}

int frb_shared_frame_pinned(FrbSharedFrame *h)
{
    Frame *f = reinterpret_cast<Frame *>(h);
    return f && f->pinned ? 1 : 0;
}

int frb_shared_frame_barrier(FrbSharedFrame *h)
{
    Frame *f = reinterpret_cast<Frame *>(h);
    if (!f) return frb_internal_fail(FRB_ERR_INVALID, "null shared frame");
    if (f->ranks == 1) return FRB_OK;
    Header *hd = f->hdr;
    const uint32_t my = f->local_sense ^ 1u;
    f->local_sense = my;
    if (hd->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == f->ranks) {
        hd->arrived.store(0, std::memory_order_relaxed);
        hd->sense.store(my, std::memory_order_release);
        return FRB_OK;
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (uint32_t spins = 0; hd->sense.load(std::memory_order_acquire) != my; ++spins) {
        if (spins < 20000u) {
            cpu_relax();
        } else {
            std::this_thread::yield();
            if ((spins & 1023u) == 0) {
                if (is_stale(f->fd, f->name))
                    return frb_internal_fail(FRB_ERR_INVALID, "shared frame %s: rank %u is attached to a frame that has been removed or replaced "
                                             "(left over from another job, or rank 0 has closed it); open it again", f->name.c_str(), f->rank);
                if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(60))
                    return frb_internal_fail(FRB_ERR_INVALID, "shared frame barrier: rank %u waited 60 s for the others", f->rank);
            }
        }
    }
    return FRB_OK;
}

void frb_shared_frame_close(FrbSharedFrame *h)
{
    Frame *f = reinterpret_cast<Frame *>(h);
    if (!f) return;
    if (f->pinned && cudaHostUnregister(f->data) != cudaSuccess) cudaGetLastError();
    if (f->rank == 0) unlink(f->name.c_str());  // the mappings of the other ranks stay valid until they unmap
    if (f->fd >= 0) close(f->fd);
    munmap(f->hdr, f->map_bytes);
    delete f;
}

}  // extern "C"
