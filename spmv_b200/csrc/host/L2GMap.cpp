// L2GMap.cpp -- halo plan construction and the NCCL-based ghost exchange.
// Reference: spmv/L2GMap.cpp:346-479 (plan, default branch), :502-642 (update
// variants), :868-959 (dispatch, reverse update).  The plan's integer artefacts are
// defined by SURVEY.md Appendix A and must equal the reference's bit for bit.
#include "L2GMap.h"

#include "b200_executor.h"
#include "device_executor.h"
#include "host_util.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

using namespace spmv;

namespace
{
const B200Executor* as_b200(const std::shared_ptr<DeviceExecutor>& e)
{
  return dynamic_cast<const B200Executor*>(e.get());
}
} // namespace

//-----------------------------------------------------------------------------
L2GMap::L2GMap(MPI_Comm comm, std::int64_t local_size, const std::vector<std::int64_t>& ghosts,
               std::shared_ptr<DeviceExecutor> exec, CommunicationModel cm)
    : _exec(exec), _ghosts(ghosts), _comm(comm), _cm(cm)
{
  if (cm != CommunicationModel::p2p_blocking && cm != CommunicationModel::p2p_nonblocking
      && cm != CommunicationModel::collective_blocking && cm != CommunicationModel::collective_nonblocking)
    throw std::runtime_error("L2GMap: one-sided / shared-memory communication models are MPI-RMA specific "
                             "and not provided by the B200 build");

  int mpi_size = 0;
  CHECK_MPI(MPI_Comm_size(comm, &mpi_size));
  CHECK_MPI(MPI_Comm_rank(comm, &_mpi_rank));

  // Step 1: ownership ranges = exclusive prefix sum of everybody's local size
  _ranges.assign(mpi_size + 1, 0);
  CHECK_MPI(MPI_Allgather(&local_size, 1, MPI_INT64_T, _ranges.data() + 1, 1, MPI_INT64_T, comm));
  for (int i = 0; i < mpi_size; ++i)
    _ranges[i + 1] += _ranges[i];
  const std::int64_t r0 = _ranges[_mpi_rank], r1 = _ranges[_mpi_rank + 1];

  if (!std::is_sorted(_ghosts.begin(), _ghosts.end()))
    throw std::runtime_error("Ghosts must be sorted");

  // Step 2: owner, owner-local index and local slot of every ghost
  std::vector<std::int32_t> ghost_count(mpi_size, 0);
  std::vector<std::int32_t> ghost_local(_ghosts.size());
  for (std::size_t k = 0; k < _ghosts.size(); ++k) {
    const std::int64_t g = _ghosts[k];
    if (g >= r0 && g < r1)
      throw std::runtime_error("Ghost index in local range");
    if (g < 0 || g >= _ranges.back())
      throw std::runtime_error("Ghost index outside the global range");
    _global_to_local.insert({g, static_cast<std::int32_t>(local_size + k)});
    const int p = static_cast<int>(std::upper_bound(_ranges.begin(), _ranges.end(), g) - _ranges.begin()) - 1;
    ++ghost_count[p];
    ghost_local[k] = static_cast<std::int32_t>(g - _ranges[p]);
  }

  // Step 3: who wants how many of my entries
  std::vector<std::int32_t> remote_count(mpi_size, 0);
  CHECK_MPI(MPI_Alltoall(ghost_count.data(), 1, MPI_INT, remote_count.data(), 1, MPI_INT, comm));

  // Step 4: symmetric neighbour list, ascending rank; zero counts kept so the
  // count arrays line up with _neighbours
  for (int q = 0; q < mpi_size; ++q) {
    if (ghost_count[q] > 0 || remote_count[q] > 0) {
      _neighbours.push_back(q);
      _send_count.push_back(ghost_count[q]);
      _recv_count.push_back(remote_count[q]);
    }
  }
  const int nnbr = static_cast<int>(_neighbours.size());
  if (nnbr == 0) {
    _send_count = {0};
    _recv_count = {0};
  }

  // Step 5: exclusive scans with the total appended
  _send_offset = {0};
  for (std::int32_t c : _send_count)
    _send_offset.push_back(_send_offset.back() + c);
  _recv_offset = {0};
  for (std::int32_t c : _recv_count)
    _recv_offset.push_back(_recv_offset.back() + c);
  _num_indices = _recv_offset.back();

  // Step 6: tell every owner which of its entries I need (owner-local indices, in my
  // ghost order); what arrives is my pack list, neighbour by neighbour
  _indexbuf_host.assign(_num_indices, 0);
  {
    MPI_Comm nbr_comm = MPI_COMM_NULL;
    CHECK_MPI(MPI_Dist_graph_create_adjacent(comm, nnbr, _neighbours.data(), MPI_UNWEIGHTED, nnbr,
                                             _neighbours.data(), MPI_UNWEIGHTED, MPI_INFO_NULL, false,
                                             &nbr_comm));
    CHECK_MPI(MPI_Neighbor_alltoallv(ghost_local.data(), _send_count.data(), _send_offset.data(),
                                     MPI_INT32_T, _indexbuf_host.data(), _recv_count.data(),
                                     _recv_offset.data(), MPI_INT32_T, nbr_comm));
    CHECK_MPI(MPI_Comm_free(&nbr_comm));
  }

  // Step 8: offsets now address the ghost tail of a vector directly
  for (std::int32_t& s : _send_offset)
    s += static_cast<std::int32_t>(local_size);

  if (_num_indices > 0) {
    _indexbuf = _exec->alloc<std::int32_t>(_num_indices);
    _exec->copy_from<std::int32_t>(_indexbuf, _exec->get_host(), _indexbuf_host.data(), _num_indices);
  }

  // The NCCL communicator is created once per executor, collectively.
  if (const B200Executor* gpu = as_b200(_exec)) {
    gpu->attach_comm(comm);
    if (mpi_size > 1)
      setup_peer_halo();
  }
}
//-----------------------------------------------------------------------------
// Collective.  Every rank exports a staging buffer and tells the others where their
// slices land in it; neighbours map each other's buffers.  If any rank cannot (ranks
// sharing a process, IPC or stream memory operations unavailable, B200_HALO=nccl) all
// ranks keep the NCCL send/recv exchange.
void L2GMap::setup_peer_halo()
{
  const B200Executor* gpu = as_b200(_exec);
  int mpi_size = 0;
  CHECK_MPI(MPI_Comm_size(_comm, &mpi_size));
  const int nnbr = static_cast<int>(_neighbours.size());
  const std::int64_t local_size = _ranges[_mpi_rank + 1] - _ranges[_mpi_rank];
  const char* mode = std::getenv("B200_HALO");
  int ok = !(mode && std::strcmp(mode, "nccl") == 0);
  detail::check(b200_set_device(gpu->get_device_id()), "cudaSetDevice");

  struct Card {
    char handle[B200_IPC_HANDLE_BYTES];
    std::int64_t slot_bytes;
    std::int64_t rev_base;
    std::int64_t rev_slot_bytes;
    std::int64_t kb_base;
    std::int64_t ok;
  };
  Card mine;
  std::memset(&mine, 0, sizeof(mine));
  _slot_bytes = (sizeof(double) * std::max<size_t>(_ghosts.size(), 1) + 255) / 256 * 256;
  // reverse_update lands in two more slots of _num_indices elements, with their own flags:
  // [fwd slot 0 | fwd slot 1 | fwd flags | pad | rev slot 0 | rev slot 1 | rev flags]
  _rev_slot_bytes = (sizeof(double) * std::max<size_t>(_num_indices, 1) + 255) / 256 * 256;
  const size_t flag_bytes = (sizeof(std::uint32_t) * static_cast<size_t>(mpi_size) + 255) / 256 * 256;
  _rev_base = 2 * _slot_bytes + flag_bytes;
  // ... and behind those the kernel-driven exchange's own channel (slots + flags): the two forward exchanges never
  // share a flag word or a slot, so each keeps its own epoch count (host-side for the copy-engine one, device-side
  // for the kernel one) and they can be used in turn on one map
  _kb_base = _rev_base + 2 * _rev_slot_bytes + flag_bytes;
  const size_t total = _kb_base + 2 * _slot_bytes + flag_bytes + 256; // + a scratch line for the probe
  void* stage = nullptr;
  if (ok && b200_ipc_alloc(&stage, total, mine.handle) != B200_OK)
    ok = 0;
  // stream memory operations present?  Probe on a scratch word past every flag (the pad behind the
  // reverse flags), and drain the stream before anybody can learn this buffer's address: a probe write
  // still in flight must not be able to clobber an epoch a fast neighbour has already published.
  if (ok
      && (b200_stream_write_u32(gpu->comm_stream(), static_cast<char*>(stage) + total - sizeof(std::uint32_t), 0u) != B200_OK
          || b200_stream_synchronize(gpu->comm_stream()) != B200_OK))
    ok = 0;
  _stage = static_cast<char*>(stage);
  mine.slot_bytes = static_cast<std::int64_t>(_slot_bytes);
  mine.rev_base = static_cast<std::int64_t>(_rev_base);
  mine.rev_slot_bytes = static_cast<std::int64_t>(_rev_slot_bytes);
  mine.kb_base = static_cast<std::int64_t>(_kb_base);
  mine.ok = ok;

  std::vector<Card> cards(mpi_size);
  CHECK_MPI(MPI_Allgather(&mine, static_cast<int>(sizeof(Card)), MPI_CHAR, cards.data(), static_cast<int>(sizeof(Card)),
                          MPI_CHAR, _comm));
  // land_mine[q]: element offset in my ghost tail where rank q's slice goes (-1: none)
  std::vector<std::int64_t> land_mine(mpi_size, -1), land_all(static_cast<size_t>(mpi_size) * mpi_size, -1);
  for (int i = 0; i < nnbr; ++i)
    land_mine[_neighbours[i]] = _send_offset[i] - local_size;
  CHECK_MPI(MPI_Allgather(land_mine.data(), mpi_size, MPI_INT64_T, land_all.data(), mpi_size, MPI_INT64_T, _comm));
  // reverse direction: rank q's ghost values that I own land at element _recv_offset[i] of my buffer
  std::vector<std::int64_t> rland_mine(mpi_size, -1), rland_all(static_cast<size_t>(mpi_size) * mpi_size, -1);
  for (int i = 0; i < nnbr; ++i)
    rland_mine[_neighbours[i]] = _recv_offset[i];
  CHECK_MPI(MPI_Allgather(rland_mine.data(), mpi_size, MPI_INT64_T, rland_all.data(), mpi_size, MPI_INT64_T, _comm));
  for (const Card& c : cards)
    ok = ok && c.ok != 0;

  _peer_stage.assign(nnbr, nullptr);
  _peer_slot_bytes.assign(nnbr, 0);
  _peer_rev_base.assign(nnbr, 0);
  _peer_rev_slot_bytes.assign(nnbr, 0);
  _peer_kb_base.assign(nnbr, 0);
  _land.assign(nnbr, -1);
  _rev_land.assign(nnbr, -1);
  if (ok) {
    for (int i = 0; i < nnbr; ++i) {
      const int q = _neighbours[i];
      void* p = nullptr;
      if (b200_ipc_open(&p, cards[q].handle) != B200_OK) {
        ok = 0;
        break;
      }
      _peer_stage[i] = static_cast<char*>(p);
      _peer_slot_bytes[i] = static_cast<size_t>(cards[q].slot_bytes);
      _peer_rev_base[i] = static_cast<size_t>(cards[q].rev_base);
      _peer_rev_slot_bytes[i] = static_cast<size_t>(cards[q].rev_slot_bytes);
      _peer_kb_base[i] = static_cast<size_t>(cards[q].kb_base);
      _land[i] = land_all[static_cast<size_t>(q) * mpi_size + _mpi_rank];
      _rev_land[i] = rland_all[static_cast<size_t>(q) * mpi_size + _mpi_rank];
    }
  }
  int all_ok = 0;
  CHECK_MPI(MPI_Allreduce(&ok, &all_ok, 1, MPI_INT, MPI_MIN, _comm));
  _peer_ok = all_ok != 0;
  // Kernel-driven exchange (push + wait kernels, device-side epoch; b200spmv.h), on its own channel of the staging
  // buffer.  It is what a captured CUDA graph can replay, so spmv::cg switches to it for the solve
  // (set_kernel_exchange); plain update() calls keep the copy-engine exchange, which costs no SM time next to the
  // persistent SpMV kernel.  B200_HALO=kernel uses it for every update, B200_HALO=copy never.
  int want_kernel = _peer_ok && nnbr <= B200_HALO_MAX_NEIGHBOURS && !(mode && std::strcmp(mode, "copy") == 0);
  if (want_kernel && nnbr >= 1) {
    std::vector<void*> ps0(nnbr), ps1(nnbr), pf(nnbr), mf(nnbr);
    for (int i = 0; i < nnbr; ++i) {
      char* pb = _peer_stage[i] + _peer_kb_base[i];
      ps0[i] = pb;
      ps1[i] = pb + _peer_slot_bytes[i];
      pf[i] = pb + 2 * _peer_slot_bytes[i] + sizeof(std::uint32_t) * _mpi_rank;
      mf[i] = _stage + _kb_base + 2 * _slot_bytes + sizeof(std::uint32_t) * _neighbours[i];
    }
    if (b200_halo_desc_create(&_halo_desc, nnbr, _recv_offset.data(), _land.data(), ps0.data(), ps1.data(), pf.data(),
                              _stage + _kb_base, _stage + _kb_base + _slot_bytes, mf.data(),
                              static_cast<std::int64_t>(_ghosts.size()), gpu->get_cuda_stream()) != B200_OK)
      want_kernel = 0;
  }
  _use_kernel = mode && std::strcmp(mode, "kernel") == 0;
  int all_kernel = 0;
  CHECK_MPI(MPI_Allreduce(&want_kernel, &all_kernel, 1, MPI_INT, MPI_MIN, _comm));
  _halo_kernel = all_kernel != 0;
  if (!_halo_kernel && _halo_desc) {
    b200_halo_desc_destroy(_halo_desc);
    _halo_desc = nullptr;
  }
  if (!_peer_ok) {
    for (char*& p : _peer_stage) {
      b200_ipc_close(p);
      p = nullptr;
    }
    if (_stage)
      b200_free(_stage);
    _stage = nullptr;
  }
  setup_peer_reduce(_peer_ok);
}
//-----------------------------------------------------------------------------
// Collective.  Mailboxes for the scalar allreduce of the CG loop (b200_peer_allreduce_sum_f64):
// every rank maps every other rank's.  Falls back to NCCL (context stays null) if any rank
// cannot, if peer memory was refused for the halo, or with B200_ALLREDUCE=nccl.
void L2GMap::setup_peer_reduce(bool peers_usable)
{
  int mpi_size = 0;
  CHECK_MPI(MPI_Comm_size(_comm, &mpi_size));
  const char* mode = std::getenv("B200_ALLREDUCE");
  int ok = peers_usable && mpi_size <= B200_PEER_REDUCE_MAX_RANKS && !(mode && std::strcmp(mode, "nccl") == 0);
  struct Card {
    char handle[B200_IPC_HANDLE_BYTES];
    std::int64_t ok;
  };
  Card mine;
  std::memset(&mine, 0, sizeof(mine));
  size_t bytes = 0;
  void* box = nullptr;
  if (ok && (b200_peer_reduce_mailbox_bytes(mpi_size, &bytes) != B200_OK || b200_ipc_alloc(&box, bytes, mine.handle) != B200_OK))
    ok = 0;
  _ar_box = static_cast<char*>(box);
  mine.ok = ok;
  std::vector<Card> cards(mpi_size);
  CHECK_MPI(MPI_Allgather(&mine, static_cast<int>(sizeof(Card)), MPI_CHAR, cards.data(), static_cast<int>(sizeof(Card)),
                          MPI_CHAR, _comm));
  for (const Card& c : cards)
    ok = ok && c.ok != 0;
  _ar_boxes.assign(mpi_size, nullptr);
  if (ok) {
    for (int q = 0; q < mpi_size && ok; ++q) {
      if (q == _mpi_rank)
        _ar_boxes[q] = _ar_box;
      else if (b200_ipc_open(&_ar_boxes[q], cards[q].handle) != B200_OK)
        ok = 0;
    }
  }
  if (ok && b200_peer_reduce_create(&_ar_ctx, _mpi_rank, mpi_size, _ar_boxes.data(), as_b200(_exec)->get_cuda_stream()) != B200_OK)
    ok = 0;
  int all_ok = 0;
  CHECK_MPI(MPI_Allreduce(&ok, &all_ok, 1, MPI_INT, MPI_MIN, _comm)); // also the barrier after the mailboxes were zeroed
  if (!all_ok)
    teardown_peer_reduce();
}
void L2GMap::allreduce_sum(double* d_val, int n, bool force_nccl) const
{
  const B200Executor* gpu = as_b200(_exec);
  if (!gpu)
    throw std::runtime_error("L2GMap::allreduce_sum needs the B200 executor");
  detail::check(b200_set_device(gpu->get_device_id()), "cudaSetDevice");
  if (_ar_ctx && !force_nccl)
    detail::check(b200_peer_allreduce_sum_f64(_ar_ctx, d_val, n, gpu->get_cuda_stream()), "peer allreduce");
  else
    detail::check(b200_allreduce_sum_f64(gpu->nccl_comm(), d_val, n, gpu->get_cuda_stream()), "nccl allreduce");
}
void L2GMap::teardown_peer_reduce()
{
  b200_peer_reduce_destroy(_ar_ctx);
  _ar_ctx = nullptr;
  for (size_t q = 0; q < _ar_boxes.size(); ++q)
    if (_ar_boxes[q] && static_cast<int>(q) != _mpi_rank)
      b200_ipc_close(_ar_boxes[q]);
  _ar_boxes.clear();
  if (_ar_box)
    b200_free(_ar_box);
  _ar_box = nullptr;
}
//-----------------------------------------------------------------------------
template <typename T>
void L2GMap::exchange_forward_peer(T* vec_data, void* stream) const
{
  int mpi_size = static_cast<int>(_ranges.size()) - 1;
  (void)mpi_size;
  const int nnbr = static_cast<int>(_neighbours.size());
  const std::uint32_t epoch = ++_epoch;
  const size_t slot = epoch & 1u;
  const std::int64_t local_size = _ranges[_mpi_rank + 1] - _ranges[_mpi_rank];
  // push my slices (L2GMap.cpp:586-598 naming: I SEND _recv_count[i] from _recv_offset[i])
  for (int i = 0; i < nnbr; ++i) {
    char* peer = _peer_stage[i];
    if (_recv_count[i] > 0)
      detail::check(b200_memcpy_d2d_async(peer + slot * _peer_slot_bytes[i] + sizeof(T) * static_cast<size_t>(_land[i]),
                                          static_cast<const char*>(_send_buf) + sizeof(T) * static_cast<size_t>(_recv_offset[i]),
                                          sizeof(T) * static_cast<size_t>(_recv_count[i]), stream),
                    "peer halo copy");
    detail::check(b200_stream_write_u32(stream, peer + 2 * _peer_slot_bytes[i] + sizeof(std::uint32_t) * _mpi_rank, epoch),
                  "peer halo flag");
  }
  // wait for every neighbour's slice of this epoch, then move the slot into the ghost tail
  for (int i = 0; i < nnbr; ++i)
    detail::check(b200_stream_wait_u32_geq(stream, _stage + 2 * _slot_bytes + sizeof(std::uint32_t) * _neighbours[i], epoch),
                  "peer halo wait");
  if (!_ghosts.empty())
    detail::check(b200_memcpy_d2d_async(vec_data + local_size, _stage + slot * _slot_bytes, sizeof(T) * _ghosts.size(),
                                        stream),
                  "peer halo unpack");
}
//-----------------------------------------------------------------------------
L2GMap::~L2GMap()
{
  // (the reference leaks _indexbuf and, for blocking models, _send_buf:
  //  L2GMap.cpp:482-499)
  try {
    _exec->synchronize();
  } catch (...) {
  }
  _exec->free(_indexbuf);
  _exec->free(_send_buf);
  _exec->free(_recv_buf);
  b200_halo_desc_destroy(_halo_desc);
  for (char* p : _peer_stage)
    b200_ipc_close(p);
  if (_stage)
    b200_free(_stage);
  teardown_peer_reduce();
}
//-----------------------------------------------------------------------------
template <typename T>
void L2GMap::ensure_send_buf() const
{
  const size_t need = sizeof(T) * static_cast<size_t>(std::max(_num_indices, 1));
  if (_send_buf == nullptr || _send_buf_bytes < need) {
    _exec->free(_send_buf);
    _send_buf = _exec->alloc<char>(need);
    _send_buf_bytes = need;
  }
}
//-----------------------------------------------------------------------------
namespace
{
int halo_push(void* d, int n, const std::int32_t* idx, const double* x, void* st) { return b200_halo_push_f64(d, n, idx, x, st); }
int halo_push(void* d, int n, const std::int32_t* idx, const float* x, void* st) { return b200_halo_push_f32(d, n, idx, x, st); }
int halo_wait(void* d, std::int64_t n, double* tail, void* st) { return b200_halo_wait_unpack_f64(d, n, tail, st); }
int halo_wait(void* d, std::int64_t n, float* tail, void* st) { return b200_halo_wait_unpack_f32(d, n, tail, st); }
} // namespace
//-----------------------------------------------------------------------------
// Kernel-driven exchange: push now; the wait + unpack follows at once (blocking models) or in
// update_finalise (overlapping models), all on the executor's stream.
template <typename T>
void L2GMap::finish_kernel_halo() const
{
  if (!_pending_tail)
    return;
  const B200Executor* gpu = as_b200(_exec);
  detail::check(halo_wait(_halo_desc, static_cast<std::int64_t>(_ghosts.size()), static_cast<T*>(_pending_tail),
                          gpu->get_cuda_stream()),
                "halo wait + unpack");
  _pending_tail = nullptr;
}
//-----------------------------------------------------------------------------
L2GMap::KernelPush L2GMap::kernel_push() const
{
  KernelPush k;
  if (_halo_kernel && _use_kernel && !_neighbours.empty()) {
    k.desc = _halo_desc;
    k.d_index = _indexbuf;
    k.n_index = _num_indices;
  }
  return k;
}
//-----------------------------------------------------------------------------
void L2GMap::adopt_pushed(double* vec_data) const
{
  if (!(_halo_kernel && _use_kernel) || _neighbours.empty())
    throw std::runtime_error("L2GMap::adopt_pushed: the kernel-driven exchange is not active");
  if (_pending_tail)
    throw std::runtime_error("L2GMap::adopt_pushed: an exchange is still waiting for its finalise");
  _pending_tail = vec_data + (_ranges[_mpi_rank + 1] - _ranges[_mpi_rank]);
  if (!overlapping())
    finish_kernel_halo<double>();
}
//-----------------------------------------------------------------------------
template <typename T>
void L2GMap::exchange_forward(T* vec_data, bool overlap) const
{
  if (_halo_kernel && _use_kernel) {
    const B200Executor* gpu = as_b200(_exec);
    if (_neighbours.empty())
      return;
    detail::check(b200_set_device(gpu->get_device_id()), "cudaSetDevice");
    finish_kernel_halo<T>(); // an exchange the caller never finalised: its slot must be drained before the next push
    // On the compute stream, in front of whatever the caller enqueues next: a push on the side stream can be
    // starved by the persistent SpMV kernel (it holds every SM's registers), and then the neighbours' wait kernels
    // stall on it -- measured: 0.28 ms per symmetric 7-point step at 8 GPUs against 0.25 with the push in line.
    detail::check(halo_push(_halo_desc, _num_indices, _indexbuf, vec_data, gpu->get_cuda_stream()), "halo push");
    _pending_tail = vec_data + (_ranges[_mpi_rank + 1] - _ranges[_mpi_rank]);
    if (!overlap)
      finish_kernel_halo<T>();
    return;
  }
  if (_neighbours.empty() && _num_indices == 0)
    return; // one rank, or a rank nobody exchanges with: nothing to pack, nothing to allocate
  ensure_send_buf<T>();
  // the previous exchange may still be reading the send buffer on the side stream if the
  // caller never finalised it
  if (const B200Executor* g = as_b200(_exec))
    if (overlap && !_neighbours.empty())
      detail::check(b200_stream_wait_event(g->get_cuda_stream(), g->arrived_event()), "stream wait");
  // pack: send_buf[k] = vec[indexbuf[k]]  (throws on the host executor: no CPU path)
  _exec->gather_ghosts_run(_num_indices, _indexbuf, vec_data, static_cast<T*>(_send_buf));

  const B200Executor* gpu = as_b200(_exec);
  if (gpu == nullptr)
    throw std::runtime_error("L2GMap::update needs the B200 executor");
  const int nnbr = static_cast<int>(_neighbours.size());
  if (nnbr == 0)
    return;
  detail::check(b200_set_device(gpu->get_device_id()), "cudaSetDevice");
  void* stream = gpu->get_cuda_stream();
  if (overlap) {
    // side stream: starts once the pack is done, signals arrival with an event
    detail::check(b200_event_record(gpu->packed_event(), gpu->get_cuda_stream()), "event record");
    detail::check(b200_stream_wait_event(gpu->comm_stream(), gpu->packed_event()), "stream wait");
    stream = gpu->comm_stream();
  }
  if (_peer_ok) {
    exchange_forward_peer(vec_data, stream);
    if (overlap)
      detail::check(b200_event_record(gpu->arrived_event(), gpu->comm_stream()), "event record");
    return;
  }
  // NB the naming (L2GMap.cpp:520-525): _recv_* describe what I SEND in a forward
  // update, _send_* what I RECEIVE (my ghosts), landing straight in the tail of vec_data.
  detail::check(b200_halo_exchange(gpu->nccl_comm(), nnbr, _neighbours.data(), _send_buf, _recv_count.data(),
                                   _recv_offset.data(), vec_data, _send_count.data(), _send_offset.data(),
                                   static_cast<int>(sizeof(T)), stream),
                "halo exchange");
  if (overlap)
    detail::check(b200_event_record(gpu->arrived_event(), gpu->comm_stream()), "event record");
}
//-----------------------------------------------------------------------------
template <typename T>
void L2GMap::update(T* vec_data) const
{
  detail::NvtxRange r("halo:pack + exchange");
  exchange_forward(vec_data, overlapping());
}
//-----------------------------------------------------------------------------
template <typename T>
void L2GMap::update_finalise(T*) const
{
  if (!overlapping() || _neighbours.empty())
    return;
  detail::NvtxRange r("halo:wait + unpack");
  const B200Executor* gpu = as_b200(_exec);
  if (gpu == nullptr)
    throw std::runtime_error("L2GMap::update_finalise needs the B200 executor");
  detail::check(b200_set_device(gpu->get_device_id()), "cudaSetDevice");
  if (_halo_kernel && _use_kernel) {
    finish_kernel_halo<T>();
    return;
  }
  detail::check(b200_stream_wait_event(gpu->get_cuda_stream(), gpu->arrived_event()), "stream wait");
}
//-----------------------------------------------------------------------------
template <typename T>
void L2GMap::reverse_update(T* vec_data) const
{
  // The reference acts only for the two blocking models and is host-only
  // (L2GMap.cpp:908-959); here every in-scope model sends the ghost tail to the
  // owners and accumulates with a scatter-add kernel, in stream order.
  const B200Executor* gpu = as_b200(_exec);
  if (gpu == nullptr)
    throw std::runtime_error("L2GMap::reverse_update needs the B200 executor");
  const int nnbr = static_cast<int>(_neighbours.size());
  if (nnbr == 0)
    return;
  const size_t need = sizeof(T) * static_cast<size_t>(std::max(_num_indices, 1));
  if (!_peer_ok && (_recv_buf == nullptr || _recv_buf_bytes < need)) {
    _exec->free(_recv_buf);
    _recv_buf = _exec->alloc<char>(need);
    _recv_buf_bytes = need;
  }
  detail::check(b200_set_device(gpu->get_device_id()), "cudaSetDevice");
  if (_peer_ok) {
    // The forward peer halo run backwards: my ghost-tail slices go straight into their owners' reverse
    // slots (copy engine over NVLink), a stream-ordered flag write publishes each, and the scatter-add
    // reads the slot once every neighbour's flag has arrived.  Same two-slot argument as the forward
    // exchange: a neighbour's push of epoch e+1 is stream-ordered after its wait for my epoch e.
    void* stream = gpu->get_cuda_stream();
    const std::uint32_t epoch = ++_rev_epoch;
    const size_t slot = epoch & 1u;
    for (int i = 0; i < nnbr; ++i) {
      char* peer = _peer_stage[i] + _peer_rev_base[i];
      if (_send_count[i] > 0)
        detail::check(b200_memcpy_d2d_async(peer + slot * _peer_rev_slot_bytes[i] + sizeof(T) * static_cast<size_t>(_rev_land[i]),
                                            vec_data + _send_offset[i], sizeof(T) * static_cast<size_t>(_send_count[i]),
                                            stream),
                      "peer reverse halo copy");
      detail::check(b200_stream_write_u32(stream, peer + 2 * _peer_rev_slot_bytes[i] + sizeof(std::uint32_t) * _mpi_rank, epoch),
                    "peer reverse halo flag");
    }
    char* mine = _stage + _rev_base;
    for (int i = 0; i < nnbr; ++i)
      detail::check(b200_stream_wait_u32_geq(stream, mine + 2 * _rev_slot_bytes + sizeof(std::uint32_t) * _neighbours[i], epoch),
                    "peer reverse halo wait");
    gpu->scatter_add_run(_num_indices, _indexbuf, reinterpret_cast<const T*>(mine + slot * _rev_slot_bytes), vec_data);
    return;
  }
  detail::check(b200_halo_exchange(gpu->nccl_comm(), nnbr, _neighbours.data(), vec_data, _send_count.data(),
                                   _send_offset.data(), _recv_buf, _recv_count.data(), _recv_offset.data(),
                                   static_cast<int>(sizeof(T)), gpu->get_cuda_stream()),
                "reverse halo exchange");
  gpu->scatter_add_run(_num_indices, _indexbuf, static_cast<const T*>(_recv_buf), vec_data);
}
//-----------------------------------------------------------------------------
std::int32_t L2GMap::global_to_local(std::int64_t i) const
{
  const std::int64_t r0 = _ranges[_mpi_rank], r1 = _ranges[_mpi_rank + 1];
  if (i >= r0 && i < r1)
    return static_cast<std::int32_t>(i - r0);
  auto it = _global_to_local.find(i);
  if (it == _global_to_local.end())
    throw std::runtime_error("L2GMap::global_to_local: index is neither owned nor a ghost");
  return it->second;
}
bool L2GMap::overlapping() const
{
  return _cm == CommunicationModel::p2p_nonblocking || _cm == CommunicationModel::collective_nonblocking;
}
std::int32_t L2GMap::local_size() const
{
  return static_cast<std::int32_t>(_ranges[_mpi_rank + 1] - _ranges[_mpi_rank]);
}
std::int32_t L2GMap::num_ghosts() const { return static_cast<std::int32_t>(_ghosts.size()); }
std::int64_t L2GMap::global_size() const { return _ranges.back(); }
std::int64_t L2GMap::global_offset() const { return _ranges[_mpi_rank]; }
//-----------------------------------------------------------------------------
template void spmv::L2GMap::update<double>(double*) const;
template void spmv::L2GMap::update<float>(float*) const;
template void spmv::L2GMap::update_finalise<double>(double*) const;
template void spmv::L2GMap::update_finalise<float>(float*) const;
template void spmv::L2GMap::reverse_update<double>(double*) const;
template void spmv::L2GMap::reverse_update<float>(float*) const;
