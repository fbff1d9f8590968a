// Node-local communicator for the z-slab decomposition: one process per GPU,
// all on one box (the 8 B200s of one node), rendezvous through a POSIX shared
// memory segment.  It carries
//   * the CUDA IPC handles of every rank's concentration buffers, so that a
//     rank can copy its boundary planes straight into its neighbours' ghost
//     planes over NVLink (cudaMemcpyAsync on peer-mapped memory);
//   * per-(rank, direction) 32-bit sequence flags that the GPUs' stream
//     front-ends write and wait on (cuStreamWriteValue32 / cuStreamWaitValue32)
//     — the halo handshake never involves the host;
//   * a host-side allreduce-max (the step-size controller's error norms, two
//     doubles per step) and a barrier, both a few shared-memory round trips.
// The reference has no distributed layer; this is the
// B200-native counterpart of section 8(e).
#pragma once

#include <atomic>
#include <cstddef>
#include <cstdint>
#include <string>

namespace smeb200 {

constexpr int kCommMaxRanks = 16;
constexpr int kCommMaxBuffers = 8; // 4 rotating state buffers + 3 relaxation buffers
constexpr int kCommMaxCompartments = 4;

struct CommRankSlot {
  // CUDA IPC memory handles (64 bytes each), per compartment and buffer
  unsigned char memHandle[kCommMaxCompartments][kCommMaxBuffers][64];
  std::atomic<uint64_t> published; // number of handle sets published
  std::atomic<uint64_t> reduceSeq; // last allreduce epoch this rank published
  double reduceVals[2][4];         // double-buffered by epoch parity
  int device;
  // ghost regions of the local arrays, in doubles, per compartment: the LOW
  // ghosts start at 0, the HIGH ghosts at ghostHiOff
  uint64_t ghostHiOff[kCommMaxCompartments];
  uint64_t ghostLoCnt[kCommMaxCompartments];
  uint64_t ghostHiCnt[kCommMaxCompartments];
  // CUDA IPC handle of the rank's in-kernel halo flag words (fused halo path)
  unsigned char flagHandle[64];
  int fusedHalo; // the rank exchanges halos inside its stage kernels
  int pad_[2];
};

struct CommSegment {
  std::atomic<uint64_t> magic;
  std::atomic<uint64_t> attached;    // ranks that mapped the segment
  std::atomic<uint64_t> barrierCount;
  uint32_t nRanks;
  // set by a rank whose run failed: every wait of the other ranks throws instead
  // of spinning until its timeout
  std::atomic<uint32_t> abort;
  // the process that created the segment: a segment whose creator is gone is a
  // leftover of a crashed run, whatever its attach count says
  int32_t creatorPid;
  uint32_t pad_[7];
  // halo flags: flag[r][0] = sequence of the last halo pushed into rank r's
  // LOW ghost plane (by rank r-1), flag[r][1] = into its HIGH ghost (by r+1).
  // Own 4 KiB page(s): registered with CUDA by every process.
  alignas(4096) uint32_t haloFlag[kCommMaxRanks][2][16];
  alignas(4096) CommRankSlot slot[kCommMaxRanks];
};

class ShmComm {
public:
  // rank 0 creates the segment, the others attach (waiting for it to appear);
  // returns once all n_ranks have attached.  Throws Error on timeout.
  ShmComm(const std::string &name, int rank, int nRanks,
          double timeoutSeconds = 120.0);
  ~ShmComm();
  ShmComm(const ShmComm &) = delete;
  ShmComm &operator=(const ShmComm &) = delete;

  int rank() const { return rank_; }
  int size() const { return nRanks_; }
  CommSegment *segment() { return seg_; }
  void barrier();
  // in-place max over ranks of n doubles; a NaN never wins against a number on
  // any rank (fmax), so every rank leaves with the same bits
  void allreduceMax(double *vals, int n);
  // in-place sum of n doubles, added in rank order (same bits on all ranks)
  void allreduceSumOrdered(double *vals, int n);
  // a rank gives up (its run threw): the waits of all ranks throw from now on
  void abort();
  bool aborted() const;
  // pointer + length of the page(s) holding the halo flags (for
  // cudaHostRegister)
  void *flagPages() { return &seg_->haloFlag[0][0][0]; }
  std::size_t flagBytes() const { return sizeof(CommSegment::haloFlag); }
  uint32_t *haloFlag(int rank, int side) {
    return &seg_->haloFlag[rank][side][0];
  }

private:
  void reduce4(double *vals, int n, bool sum);
  void waitUntil(const std::atomic<uint64_t> &v, uint64_t target,
                 const char *what) const;
  std::string name_;
  int rank_{0}, nRanks_{1};
  double timeout_{120.0};
  CommSegment *seg_{nullptr};
  uint64_t barrierEpoch_{0};
  uint64_t reduceEpoch_{0};
  bool owner_{false};
};

} // namespace smeb200
