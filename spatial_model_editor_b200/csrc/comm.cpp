#include "comm.hpp"
#include "cuda_api.hpp"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <cerrno>
#include <csignal>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

namespace smeb200 {

namespace {
constexpr uint64_t kMagic = 0x736d656232303063ull; // "smeb200c"

double now() {
  return std::chrono::duration<double>(
             std::chrono::steady_clock::now().time_since_epoch())
      .count();
}
} // namespace

ShmComm::ShmComm(const std::string &name, int rank, int nRanks,
                 double timeoutSeconds)
    : name_{name.empty() || name[0] == '/' ? name : "/" + name}, rank_{rank},
      nRanks_{nRanks}, timeout_{timeoutSeconds} {
  if (nRanks < 1 || nRanks > kCommMaxRanks || rank < 0 || rank >= nRanks) {
    throw Error("smeb200: invalid rank / number of ranks for the slab "
                "communicator");
  }
  if (name_.size() < 2) {
    throw Error("smeb200: the slab communicator needs a name");
  }
  int fd = -1;
  const double t0 = now();
  if (rank == 0) {
    shm_unlink(name_.c_str()); // stale segment of a crashed run
    fd = shm_open(name_.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) {
      throw Error("smeb200: shm_open(create) failed for " + name_);
    }
    if (ftruncate(fd, sizeof(CommSegment)) != 0) {
      close(fd);
      throw Error("smeb200: ftruncate failed for " + name_);
    }
    owner_ = true;
  }
  // ranks other than 0 attach to the segment rank 0 creates.  A segment of that
  // name may already exist: left behind by a crashed run, or one whose rank 0 has
  // not unlinked it yet (back-to-back simulators).  Such a segment is recognised
  // by its attach count — every rank attaches exactly once, so a count above
  // nRanks means somebody else's ranks are (or were) on it — and left alone.
  for (;;) {
    ino_t inode = 0;
    if (rank != 0) {
      for (;;) {
        fd = shm_open(name_.c_str(), O_RDWR, 0600);
        if (fd >= 0) {
          struct stat st {};
          if (fstat(fd, &st) == 0 &&
              static_cast<std::size_t>(st.st_size) >= sizeof(CommSegment)) {
            inode = st.st_ino;
            break;
          }
          close(fd);
          fd = -1;
        }
        if (now() - t0 > timeout_) {
          throw Error("smeb200: timed out waiting for rank 0 to create " + name_);
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(2));
      }
    }
    void *p = mmap(nullptr, sizeof(CommSegment), PROT_READ | PROT_WRITE,
                   MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
      throw Error("smeb200: mmap failed for " + name_);
    }
    seg_ = static_cast<CommSegment *>(p);
    if (rank == 0) {
      // a fresh segment is zero-filled; publish it
      seg_->nRanks = static_cast<uint32_t>(nRanks);
      seg_->creatorPid = static_cast<int32_t>(getpid());
      seg_->magic.store(kMagic, std::memory_order_release);
      seg_->attached.fetch_add(1, std::memory_order_acq_rel);
      break;
    }
    // the name still leads to the segment mapped here?  (rank 0 clears the magic
    // and unlinks when it is done; a new simulation creates a new segment)
    auto nameMovedOn = [&]() {
      const int probe = shm_open(name_.c_str(), O_RDWR, 0600);
      if (probe < 0) {
        return true;
      }
      struct stat st {};
      const bool moved = fstat(probe, &st) != 0 || st.st_ino != inode;
      close(probe);
      return moved;
    };
    bool stale = false;
    unsigned polls = 0;
    while (seg_->magic.load(std::memory_order_acquire) != kMagic) {
      if ((++polls & 15u) == 0 && nameMovedOn()) {
        stale = true;
        break;
      }
      if (now() - t0 > timeout_) {
        munmap(seg_, sizeof(CommSegment));
        seg_ = nullptr;
        throw Error("smeb200: timed out waiting for the communicator header");
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    if (!stale && seg_->creatorPid > 0 && kill(seg_->creatorPid, 0) != 0 &&
        errno == ESRCH) {
      stale = true; // left behind by a process that no longer exists
    }
    if (!stale && seg_->nRanks != static_cast<uint32_t>(nRanks)) {
      munmap(seg_, sizeof(CommSegment));
      seg_ = nullptr;
      throw Error("smeb200: ranks disagree about the number of ranks");
    }
    if (!stale &&
        seg_->attached.fetch_add(1, std::memory_order_acq_rel) + 1 >
            static_cast<uint64_t>(nRanks)) {
      stale = true; // a full house already: not this simulation's segment
    }
    if (!stale) {
      break;
    }
    munmap(seg_, sizeof(CommSegment));
    seg_ = nullptr;
    if (now() - t0 > timeout_) {
      throw Error("smeb200: timed out waiting for rank 0 to create " + name_);
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
  }
  waitUntil(seg_->attached, static_cast<uint64_t>(nRanks), "all ranks to attach");
}

ShmComm::~ShmComm() {
  if (seg_ != nullptr) {
    if (owner_) { // nobody may attach to a segment that is going away
      seg_->magic.store(0, std::memory_order_release);
    }
    munmap(seg_, sizeof(CommSegment));
  }
  if (owner_) {
    shm_unlink(name_.c_str());
  }
}

void ShmComm::waitUntil(const std::atomic<uint64_t> &v, uint64_t target,
                        const char *what) const {
  const double t0 = now();
  unsigned spins = 0;
  while (v.load(std::memory_order_acquire) < target) {
    if (++spins > 2000) {
      std::this_thread::yield();
      if ((spins & 0xff) == 0 &&
          seg_->abort.load(std::memory_order_acquire) != 0) {
        throw Error(std::string("smeb200: another rank of the decomposed "
                                "simulation failed (while waiting for ") +
                    what + ")");
      }
      if ((spins & 0xfff) == 0 && now() - t0 > timeout_) {
        throw Error(std::string("smeb200: timed out waiting for ") + what);
      }
    }
  }
}

void ShmComm::abort() { seg_->abort.store(1, std::memory_order_release); }

bool ShmComm::aborted() const {
  return seg_->abort.load(std::memory_order_acquire) != 0;
}

void ShmComm::barrier() {
  ++barrierEpoch_;
  seg_->barrierCount.fetch_add(1, std::memory_order_acq_rel);
  waitUntil(seg_->barrierCount, barrierEpoch_ * static_cast<uint64_t>(nRanks_),
            "the other ranks at a barrier");
}

// one round over at most 4 doubles: every rank publishes its values under the
// next epoch and folds all ranks' values in rank order, its own included, so all
// ranks compute the same bits (sum: left to right from zero; max: fmax, where a
// NaN never wins against a number)
void ShmComm::reduce4(double *vals, int n, bool sum) {
  ++reduceEpoch_;
  auto &mine = seg_->slot[rank_];
  std::memcpy(mine.reduceVals[reduceEpoch_ & 1], vals, n * sizeof(double));
  mine.reduceSeq.store(reduceEpoch_, std::memory_order_release);
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int r = 0; r < nRanks_; ++r) {
    if (r != rank_) {
      waitUntil(seg_->slot[r].reduceSeq, reduceEpoch_,
                "the other ranks' values of a reduction");
    }
    const double *v = seg_->slot[r].reduceVals[reduceEpoch_ & 1];
    for (int i = 0; i < n; ++i) {
      acc[i] = sum ? acc[i] + v[i] : (r == 0 ? v[i] : std::fmax(acc[i], v[i]));
    }
  }
  std::memcpy(vals, acc, n * sizeof(double));
}

void ShmComm::allreduceSumOrdered(double *vals, int n) {
  for (int i = 0; i < n; i += 4) {
    reduce4(vals + i, std::min(4, n - i), true);
  }
}

void ShmComm::allreduceMax(double *vals, int n) {
  for (int i = 0; i < n; i += 4) {
    reduce4(vals + i, std::min(4, n - i), false);
  }
}

} // namespace smeb200
