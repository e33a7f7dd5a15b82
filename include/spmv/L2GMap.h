// L2GMap.h -- local-to-global map and halo plan (reference: spmv/L2GMap.h:24-196).
//
// A rank owns a contiguous range of global indices; indices owned elsewhere are
// "ghosts" stored after the owned entries of every vector ([owned | ghosts ascending]).
// The constructor computes the exchange plan exactly as the reference's default
// branch does (L2GMap.cpp:346-479; semantics in SURVEY.md Appendix A) -- those
// integer artefacts are bit-identical to the reference's.  update() moves ghost
// values with a pack kernel + grouped NCCL send/recv instead of MPI.
#pragma once

#include "config.h"
#include "spmv_export.h"

#include "mpi_utils.h"

#include <cstdint>
#include <map>
#include <memory>
#include <mpi.h>
#include <vector>

namespace spmv
{

class DeviceExecutor;

class SPMV_EXPORT L2GMap
{
public:
  /// @param comm        communicator (collective call)
  /// @param local_size  number of owned entries
  /// @param ghosts      global indices owned by other ranks, sorted ascending
  /// @param exec        executor that holds the pack index list and send buffer
  /// @param cm          communication model (see mpi_utils.h)
  /// Throws std::runtime_error for unsorted ghosts, a ghost inside the local range
  /// (L2GMap.cpp:362-363,371-372) or an out-of-scope model.
  L2GMap(MPI_Comm comm, std::int64_t local_size, const std::vector<std::int64_t>& ghosts,
         std::shared_ptr<spmv::DeviceExecutor> exec,
         CommunicationModel cm = CommunicationModel::collective_blocking);
  ~L2GMap();
  L2GMap(const L2GMap&) = delete;
  L2GMap& operator=(const L2GMap&) = delete;

  std::int32_t local_size() const;
  std::int32_t num_ghosts() const;
  std::int64_t global_size() const;
  std::int64_t global_offset() const;
  std::int32_t global_to_local(std::int64_t i) const;
  bool overlapping() const;
  MPI_Comm global_comm() const { return _comm; }
  int rank() const { return _mpi_rank; }
  const std::vector<std::int64_t>& ghosts() const { return _ghosts; }

  /// Forward halo: owners' values -> ghost tail of vec_data (executor memory with
  /// room for local_size()+num_ghosts()).  Blocking models: complete in stream order
  /// on the executor's stream.  Non-blocking models: started on the side stream;
  /// update_finalise() makes the executor's stream wait for the arrival.
  template <typename T>
  void update(T* vec_data) const;
  template <typename T>
  void update_finalise(T* vec_data) const;
  /// Reverse halo: ghost values -> added onto their owners' entries
  /// (L2GMap.cpp:908-959), exchange + scatter-add kernel.
  template <typename T>
  void reverse_update(T* vec_data) const;

  // ---- the plan, for inspection (private in the reference; exposed read-only so
  //      parity of the integer artefacts can be tested) ----
  const std::vector<std::int64_t>& ranges() const { return _ranges; }
  const std::vector<int>& neighbours() const { return _neighbours; }
  /// what this rank RECEIVES in update() per neighbour (its ghosts), and where:
  /// offsets already shifted by local_size (L2GMap.cpp:460-461)
  const std::vector<std::int32_t>& send_count() const { return _send_count; }
  const std::vector<std::int32_t>& send_offset() const { return _send_offset; }
  /// what this rank SENDS in update() per neighbour (slices of the packed buffer)
  const std::vector<std::int32_t>& recv_count() const { return _recv_count; }
  const std::vector<std::int32_t>& recv_offset() const { return _recv_offset; }
  /// Context for b200_peer_allreduce_sum_f64 over this map's communicator (nullptr: one rank,
  /// or peer memory unavailable -- use the NCCL allreduce)
  void* peer_reduce() const { return _ar_ctx; }
  /// Sum n <= 4 device doubles over the communicator, in place, on the executor's stream
  /// (peer-memory kernel if set up and !force_nccl, else ncclAllReduce).  Collective.
  void allreduce_sum(double* d_val, int n, bool force_nccl = false) const;
  /// owner-local indices to pack, neighbour by neighbour (host copy of _indexbuf)
  const std::vector<std::int32_t>& indexbuf() const { return _indexbuf_host; }
  CommunicationModel communication_model() const { return _cm; }
  /// "kernel" (push + wait kernels over peer memory), "copy" (copy engine + stream memory operations),
  /// "nccl" (grouped send/recv) or "none" (one rank): how update() moves the halo
  const char* halo_mode() const
  {
    return _ranges.size() <= 2 ? "none" : (_halo_kernel && _use_kernel) ? "kernel" : _peer_ok ? "copy" : "nccl";
  }
  /// ... and the one spmv::cg switches to for its solve
  const char* halo_mode_cg() const { return (_ranges.size() > 2 && _halo_kernel) ? "kernel" : halo_mode(); }
  /// Switch update() between the copy-engine exchange (default: no SM time next to the SpMV kernel) and the
  /// kernel-driven one (device-side epochs: what a captured CUDA graph can replay); returns the previous setting.
  /// No-op (false) where the kernel exchange was not set up.  Every rank must switch at the same point, with no
  /// update in flight.
  bool set_kernel_exchange(bool on) const
  {
    const bool was = _use_kernel;
    if (_halo_kernel)
      _use_kernel = on;
    return was;
  }

  /// For a kernel that performs the PUSH half of the kernel-driven exchange itself (the fused CG step,
  /// b200_cg_fused_step_halo_f64): the descriptor, the index list and its length; desc == nullptr unless the
  /// kernel exchange is the active one and this rank has neighbours.
  struct KernelPush {
    void* desc = nullptr;
    const std::int32_t* d_index = nullptr;
    std::int32_t n_index = 0;
  };
  KernelPush kernel_push() const;
  /// The counterpart of update(vec) after such a kernel was enqueued on the executor's stream: the exchange is
  /// in flight, its wait + unpack follows at once (blocking models) or in update_finalise (overlapping models).
  void adopt_pushed(double* vec_data) const;
  /// For a kernel that performs the WAIT half itself (b200_csr_spmv_rowlist_dot_halo_f64 reads the ghost values
  /// from the staging slot): the descriptor of the exchange in flight, or nullptr if none is (blocking models,
  /// other exchange modes).  finalised_by_kernel() then stands in for update_finalise(): the ghost tail of the
  /// vector is NOT filled.
  void* kernel_wait_desc() const { return (_halo_kernel && _use_kernel && _pending_tail) ? _halo_desc : nullptr; }
  void finalised_by_kernel() const { _pending_tail = nullptr; }

private:
  template <typename T>
  void exchange_forward(T* vec_data, bool overlap) const;
  template <typename T>
  void ensure_send_buf() const;

  std::shared_ptr<spmv::DeviceExecutor> _exec;
  std::vector<std::int64_t> _ranges;
  std::int32_t _mpi_rank = -1;
  std::map<std::int64_t, std::int32_t> _global_to_local;
  std::vector<std::int64_t> _ghosts;

  int _num_indices = 0;
  std::int32_t* _indexbuf = nullptr; // executor memory
  std::vector<std::int32_t> _indexbuf_host;
  std::vector<std::int32_t> _send_count, _recv_count, _send_offset, _recv_offset;
  std::vector<int> _neighbours;

  MPI_Comm _comm = MPI_COMM_NULL;
  CommunicationModel _cm = CommunicationModel::collective_blocking;
  mutable void* _send_buf = nullptr; // executor memory, lazily allocated (L2GMap.cpp:507-514)
  mutable size_t _send_buf_bytes = 0;
  mutable void* _recv_buf = nullptr; // reverse_update staging
  mutable size_t _recv_buf_bytes = 0;

  // ---- direct peer halo (b200spmv.h "direct peer halo") ----------------
  // My staging buffer: [slot 0 | slot 1 | arrival flags, one uint32 per rank]; neighbours
  // push their slices into the slot of the current epoch and then write the epoch into
  // their flag; I wait for every neighbour's flag and copy the slot into the ghost tail.
  // Double buffering is enough: a neighbour can only be one epoch ahead of me, because
  // its next push is stream-ordered after its wait for my previous one.
  void setup_peer_halo();
  void setup_peer_reduce(bool peers_usable);
  void teardown_peer_reduce();
  template <typename T>
  void exchange_forward_peer(T* vec_data, void* stream) const;
  template <typename T>
  void finish_kernel_halo() const;
  // kernel-driven exchange (b200_halo_push / b200_halo_wait_unpack): descriptor in device memory, and the
  // ghost tail of an update whose wait + unpack is still to be enqueued (overlapping models)
  bool _halo_kernel = false;
  void* _halo_desc = nullptr;
  mutable void* _pending_tail = nullptr;
  mutable bool _use_kernel = false; // update() takes the kernel channel (B200_HALO=kernel, or spmv::cg for its solve)
  size_t _kb_base = 0;              // the kernel channel's [slot 0 | slot 1 | flags] inside the staging buffer
  std::vector<size_t> _peer_kb_base;
  bool _peer_ok = false;
  char* _stage = nullptr;            // my staging buffer (cudaMalloc, IPC-exported)
  size_t _slot_bytes = 0;
  std::vector<char*> _peer_stage;    // per neighbour: its staging buffer mapped here
  std::vector<size_t> _peer_slot_bytes;
  std::vector<std::int64_t> _land;   // per neighbour: element offset of my slice in its ghost tail
  mutable std::uint32_t _epoch = 0;
  // reverse direction (reverse_update): two more slots of _num_indices elements + flags behind the forward ones
  size_t _rev_base = 0, _rev_slot_bytes = 0;
  std::vector<size_t> _peer_rev_base, _peer_rev_slot_bytes;
  std::vector<std::int64_t> _rev_land; // per neighbour: element offset of my slice in its reverse slot
  mutable std::uint32_t _rev_epoch = 0;
  // scalar allreduce over peer memory (CG): my mailbox, everyone's as mapped here, device descriptor
  char* _ar_box = nullptr;
  std::vector<void*> _ar_boxes;
  void* _ar_ctx = nullptr;
};

} // namespace spmv
