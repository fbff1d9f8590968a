#include "comm.hpp"
#include "cuda_api.hpp"
#include <algorithm>
#include <chrono>
#include <cstring>
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
  } else {
    for (;;) {
      fd = shm_open(name_.c_str(), O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st {};
        if (fstat(fd, &st) == 0 &&
            static_cast<std::size_t>(st.st_size) >= sizeof(CommSegment)) {
          break;
        }
        close(fd);
        fd = -1;
      }
      if (now() - t0 > timeout_) {
        throw Error("smeb200: timed out waiting for rank 0 to create " +
                    name_);
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
    seg_->magic.store(kMagic, std::memory_order_release);
  } else {
    while (seg_->magic.load(std::memory_order_acquire) != kMagic) {
      if (now() - t0 > timeout_) {
        throw Error("smeb200: timed out waiting for the communicator header");
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    if (seg_->nRanks != static_cast<uint32_t>(nRanks)) {
      throw Error("smeb200: ranks disagree about the number of ranks");
    }
  }
  seg_->attached.fetch_add(1, std::memory_order_acq_rel);
  waitUntil(seg_->attached, static_cast<uint64_t>(nRanks), "all ranks to attach");
}

ShmComm::~ShmComm() {
  if (seg_ != nullptr) {
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
      if ((spins & 0xfff) == 0 && now() - t0 > timeout_) {
        throw Error(std::string("smeb200: timed out waiting for ") + what);
      }
    }
  }
}

void ShmComm::barrier() {
  ++barrierEpoch_;
  seg_->barrierCount.fetch_add(1, std::memory_order_acq_rel);
  waitUntil(seg_->barrierCount, barrierEpoch_ * static_cast<uint64_t>(nRanks_),
            "the other ranks at a barrier");
}

// in-place sum over ranks of n <= 4 doubles, added in rank order starting from
// zero on every rank: all ranks get the same bits
void ShmComm::allreduceSumOrdered(double *vals, int n) {
  n = std::min(n, 4);
  ++reduceEpoch_;
  auto &mine = seg_->slot[rank_];
  std::memcpy(mine.reduceVals[reduceEpoch_ & 1], vals, n * sizeof(double));
  mine.reduceSeq.store(reduceEpoch_, std::memory_order_release);
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int r = 0; r < nRanks_; ++r) {
    if (r != rank_) {
      waitUntil(seg_->slot[r].reduceSeq, reduceEpoch_,
                "the other ranks' partial sums");
    }
    const double *v = seg_->slot[r].reduceVals[reduceEpoch_ & 1];
    for (int i = 0; i < n; ++i) {
      acc[i] += v[i];
    }
  }
  std::memcpy(vals, acc, n * sizeof(double));
}

void ShmComm::allreduceMax(double *vals, int n) {
  n = std::min(n, 4);
  ++reduceEpoch_;
  auto &mine = seg_->slot[rank_];
  std::memcpy(mine.reduceVals[reduceEpoch_ & 1], vals, n * sizeof(double));
  mine.reduceSeq.store(reduceEpoch_, std::memory_order_release);
  for (int r = 0; r < nRanks_; ++r) {
    if (r == rank_) {
      continue;
    }
    waitUntil(seg_->slot[r].reduceSeq, reduceEpoch_,
              "the other ranks' error norms");
    const double *other = seg_->slot[r].reduceVals[reduceEpoch_ & 1];
    for (int i = 0; i < n; ++i) {
      vals[i] = std::max(vals[i], other[i]); // NaN in other[i] is not taken
    }
  }
}

} // namespace smeb200
