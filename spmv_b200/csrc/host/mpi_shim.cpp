// mpi_shim.cpp -- implementation of include/spmv/mpi_shim/mpi.h (see there).
#include <mpi.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace
{

[[noreturn]] void die(const char* what)
{
  std::fprintf(stderr, "spmv mpi shim: %s\n", what);
  std::abort();
}

// ---------------------------------------------------------------------------
// Transport interface: a barrier and a byte-wise alltoallv; optional p2p.
// ---------------------------------------------------------------------------
struct World {
  int size = 1;
  virtual ~World() = default;
  virtual void barrier(int rank) = 0;
  virtual void alltoallv(int rank, const void* sbuf, const int64_t* scnt, const int64_t* sdsp, void* rbuf,
                         const int64_t* rcnt, const int64_t* rdsp)
      = 0;
  virtual void send(int, int, int, const void*, size_t) { die("point-to-point needs the threads transport"); }
  virtual void recv(int, int, int, void*, size_t) { die("point-to-point needs the threads transport"); }
};

struct SelfWorld final : World {
  void barrier(int) override {}
  void alltoallv(int, const void* sbuf, const int64_t* scnt, const int64_t* sdsp, void* rbuf,
                 const int64_t* rcnt, const int64_t* rdsp) override
  {
    if (scnt[0] != rcnt[0])
      die("self alltoallv: count mismatch");
    if (scnt[0] > 0)
      std::memmove(static_cast<char*>(rbuf) + rdsp[0], static_cast<const char*>(sbuf) + sdsp[0], scnt[0]);
  }
  // a rank may message itself (eager copy through a queue)
  struct Msg {
    int tag;
    std::vector<char> data;
  };
  std::deque<Msg> box;
  void send(int, int, int tag, const void* buf, size_t n) override
  {
    box.push_back({tag, std::vector<char>(static_cast<const char*>(buf), static_cast<const char*>(buf) + n)});
  }
  void recv(int, int, int tag, void* buf, size_t n) override
  {
    for (auto it = box.begin(); it != box.end(); ++it)
      if (it->tag == tag) {
        if (it->data.size() != n)
          die("self recv: size mismatch");
        std::memcpy(buf, it->data.data(), n);
        box.erase(it);
        return;
      }
    die("self recv: no matching message (would deadlock)");
  }
};

struct ThreadWorld final : World {
  struct Slot {
    const void* sbuf;
    const int64_t* scnt;
    const int64_t* sdsp;
  };
  std::vector<Slot> slots;
  std::mutex mu;
  std::condition_variable cv;
  int arrived = 0;
  long generation = 0;
  struct Msg {
    int src, tag;
    std::vector<char> data;
  };
  std::vector<std::deque<Msg>> boxes;
  std::mutex box_mu;
  std::condition_variable box_cv;

  explicit ThreadWorld(int n) : slots(n), boxes(n) { size = n; }

  void barrier(int) override
  {
    std::unique_lock<std::mutex> lk(mu);
    const long gen = generation;
    if (++arrived == size) {
      arrived = 0;
      ++generation;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return generation != gen; });
    }
  }
  void alltoallv(int rank, const void* sbuf, const int64_t* scnt, const int64_t* sdsp, void* rbuf,
                 const int64_t* rcnt, const int64_t* rdsp) override
  {
    slots[rank] = {sbuf, scnt, sdsp};
    barrier(rank);
    for (int src = 0; src < size; ++src) {
      const Slot& s = slots[src];
      if (s.scnt[rank] != rcnt[src])
        die("alltoallv: send/recv count mismatch between ranks");
      if (rcnt[src] > 0)
        std::memcpy(static_cast<char*>(rbuf) + rdsp[src], static_cast<const char*>(s.sbuf) + s.sdsp[rank],
                    rcnt[src]);
    }
    barrier(rank);
  }
  void send(int rank, int dest, int tag, const void* buf, size_t n) override
  {
    {
      std::lock_guard<std::mutex> lk(box_mu);
      boxes[dest].push_back(
          {rank, tag, std::vector<char>(static_cast<const char*>(buf), static_cast<const char*>(buf) + n)});
    }
    box_cv.notify_all();
  }
  void recv(int rank, int source, int tag, void* buf, size_t n) override
  {
    std::unique_lock<std::mutex> lk(box_mu);
    for (;;) {
      auto& q = boxes[rank];
      for (auto it = q.begin(); it != q.end(); ++it)
        if (it->src == source && it->tag == tag) {
          if (it->data.size() != n)
            die("recv: message size mismatch");
          std::memcpy(buf, it->data.data(), n);
          q.erase(it);
          return;
        }
      box_cv.wait(lk);
    }
  }
};

struct CallbackWorld final : World {
  spmv_shim_barrier_fn barrier_fn;
  spmv_shim_alltoallv_fn a2a_fn;
  void* ctx;
  void barrier(int) override
  {
    if (barrier_fn(ctx) != 0)
      die("barrier callback failed");
  }
  void alltoallv(int, const void* sbuf, const int64_t* scnt, const int64_t* sdsp, void* rbuf,
                 const int64_t* rcnt, const int64_t* rdsp) override
  {
    if (a2a_fn(ctx, sbuf, scnt, sdsp, rbuf, rcnt, rdsp) != 0)
      die("alltoallv callback failed");
  }
};

} // namespace

struct spmv_shim_comm {
  std::shared_ptr<World> world;
  int rank = 0;
  bool is_graph = false;
  std::vector<int> sources, dests;
};

struct spmv_shim_request {
  // a pending receive (sends are eager and complete at once)
  spmv_shim_comm* comm = nullptr;
  void* buf = nullptr;
  size_t bytes = 0;
  int source = 0, tag = 0;
  bool is_recv = false;
};

namespace
{

spmv_shim_comm* make_self()
{
  auto* c = new spmv_shim_comm;
  c->world = std::make_shared<SelfWorld>();
  c->rank = 0;
  return c;
}

std::mutex g_world_mu;
spmv_shim_comm* g_process_world = nullptr;         // self or callbacks
thread_local spmv_shim_comm* t_thread_world = nullptr; // inside spmv_shim_run_ranks

int type_size(MPI_Datatype t)
{
  switch (t) {
  case MPI_CHAR:
  case MPI_BYTE:
    return 1;
  case MPI_INT:
  case MPI_INT32_T:
  case MPI_FLOAT:
    return 4;
  case MPI_INT64_T:
  case MPI_DOUBLE:
  case MPI_C_FLOAT_COMPLEX:
    return 8;
  case MPI_DOUBLE_COMPLEX:
    return 16;
  default:
    die("unknown MPI_Datatype");
  }
}

template <typename T>
void fold(T* acc, const T* in, int n, MPI_Op op)
{
  for (int i = 0; i < n; ++i) {
    if (op == MPI_SUM)
      acc[i] = acc[i] + in[i];
    else if (op == MPI_MIN)
      acc[i] = std::min(acc[i], in[i]);
    else if (op == MPI_MAX)
      acc[i] = std::max(acc[i], in[i]);
    else
      die("unknown MPI_Op");
  }
}

void fold_any(void* acc, const void* in, int n, MPI_Datatype t, MPI_Op op)
{
  switch (t) {
  case MPI_INT:
  case MPI_INT32_T:
    return fold(static_cast<int32_t*>(acc), static_cast<const int32_t*>(in), n, op);
  case MPI_INT64_T:
    return fold(static_cast<int64_t*>(acc), static_cast<const int64_t*>(in), n, op);
  case MPI_FLOAT:
    return fold(static_cast<float*>(acc), static_cast<const float*>(in), n, op);
  case MPI_DOUBLE:
    return fold(static_cast<double*>(acc), static_cast<const double*>(in), n, op);
  case MPI_CHAR:
  case MPI_BYTE:
    return fold(static_cast<char*>(acc), static_cast<const char*>(in), n, op);
  default:
    die("reduction on unsupported datatype");
  }
}

// full-size byte alltoallv from per-peer element counts
void a2av(spmv_shim_comm* c, const void* sbuf, const std::vector<int64_t>& scnt,
          const std::vector<int64_t>& sdsp, void* rbuf, const std::vector<int64_t>& rcnt,
          const std::vector<int64_t>& rdsp)
{
  c->world->alltoallv(c->rank, sbuf, scnt.data(), sdsp.data(), rbuf, rcnt.data(), rdsp.data());
}

} // namespace

extern "C" {

MPI_Comm spmv_shim_comm_world(void)
{
  if (t_thread_world)
    return t_thread_world;
  std::lock_guard<std::mutex> lk(g_world_mu);
  if (!g_process_world)
    g_process_world = make_self();
  return g_process_world;
}

int spmv_shim_type_size(MPI_Datatype t) { return type_size(t); }

int spmv_shim_run_ranks(int n, void (*fn)(void*), void* arg)
{
  if (n < 1)
    return 1;
  auto world = std::make_shared<ThreadWorld>(n);
  std::vector<std::thread> threads;
  std::vector<int> failed(n, 0);
  for (int r = 0; r < n; ++r)
    threads.emplace_back([&, r] {
      spmv_shim_comm c;
      c.world = world;
      c.rank = r;
      t_thread_world = &c;
      try {
        fn(arg);
      } catch (const std::exception& e) {
        std::fprintf(stderr, "spmv mpi shim: rank %d threw: %s\n", r, e.what());
        failed[r] = 1;
      } catch (...) {
        failed[r] = 1;
      }
      t_thread_world = nullptr;
    });
  for (auto& t : threads)
    t.join();
  return std::any_of(failed.begin(), failed.end(), [](int f) { return f != 0; }) ? 1 : 0;
}

int spmv_shim_set_world_callbacks(int rank, int size, spmv_shim_barrier_fn barrier,
                                  spmv_shim_alltoallv_fn alltoallv, void* ctx)
{
  if (size < 1 || rank < 0 || rank >= size)
    return MPI_ERR_OTHER;
  std::lock_guard<std::mutex> lk(g_world_mu);
  spmv_shim_comm* c = nullptr;
  if (barrier == nullptr || alltoallv == nullptr) {
    if (size != 1)
      return MPI_ERR_OTHER;
    c = make_self();
  } else {
    auto w = std::make_shared<CallbackWorld>();
    w->size = size;
    w->barrier_fn = barrier;
    w->a2a_fn = alltoallv;
    w->ctx = ctx;
    c = new spmv_shim_comm;
    c->world = w;
    c->rank = rank;
  }
  // the previous world is leaked on purpose: objects built on it may still hold it
  g_process_world = c;
  return MPI_SUCCESS;
}

int MPI_Init(int*, char***) { return MPI_SUCCESS; }
int MPI_Init_thread(int*, char***, int required, int* provided)
{
  if (provided)
    *provided = required;
  return MPI_SUCCESS;
}
int MPI_Finalize(void) { return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm comm, int* size)
{
  *size = comm->world->size;
  return MPI_SUCCESS;
}
int MPI_Comm_rank(MPI_Comm comm, int* rank)
{
  *rank = comm->rank;
  return MPI_SUCCESS;
}
int MPI_Comm_free(MPI_Comm* comm)
{
  if (comm && *comm && (*comm)->is_graph)
    delete *comm;
  if (comm)
    *comm = MPI_COMM_NULL;
  return MPI_SUCCESS;
}
int MPI_Barrier(MPI_Comm comm)
{
  comm->world->barrier(comm->rank);
  return MPI_SUCCESS;
}
double MPI_Wtime(void)
{
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}
int MPI_Pcontrol(const int, ...) { return MPI_SUCCESS; }
int MPI_Error_string(int errorcode, char* string, int* resultlen)
{
  int n = std::snprintf(string, MPI_MAX_ERROR_STRING, "spmv mpi shim error %d", errorcode);
  if (resultlen)
    *resultlen = n;
  return MPI_SUCCESS;
}

int MPI_Alltoallv(const void* sendbuf, const int* sendcounts, const int* sdispls, MPI_Datatype sendtype,
                  void* recvbuf, const int* recvcounts, const int* rdispls, MPI_Datatype recvtype,
                  MPI_Comm comm)
{
  const int n = comm->world->size;
  const int64_t ss = type_size(sendtype), rs = type_size(recvtype);
  std::vector<int64_t> sc(n), sd(n), rc(n), rd(n);
  for (int i = 0; i < n; ++i) {
    sc[i] = sendcounts[i] * ss;
    sd[i] = sdispls[i] * ss;
    rc[i] = recvcounts[i] * rs;
    rd[i] = rdispls[i] * rs;
  }
  a2av(comm, sendbuf, sc, sd, recvbuf, rc, rd);
  return MPI_SUCCESS;
}

int MPI_Alltoall(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
                 MPI_Datatype recvtype, MPI_Comm comm)
{
  const int n = comm->world->size;
  const int64_t sb = (int64_t)sendcount * type_size(sendtype), rb = (int64_t)recvcount * type_size(recvtype);
  std::vector<int64_t> sc(n, sb), sd(n), rc(n, rb), rd(n);
  for (int i = 0; i < n; ++i) {
    sd[i] = i * sb;
    rd[i] = i * rb;
  }
  a2av(comm, sendbuf, sc, sd, recvbuf, rc, rd);
  return MPI_SUCCESS;
}

int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
                  MPI_Datatype recvtype, MPI_Comm comm)
{
  const int n = comm->world->size;
  const int64_t sb = (int64_t)sendcount * type_size(sendtype), rb = (int64_t)recvcount * type_size(recvtype);
  std::vector<int64_t> sc(n, sb), sd(n, 0), rc(n, rb), rd(n);
  for (int i = 0; i < n; ++i)
    rd[i] = i * rb;
  a2av(comm, sendbuf, sc, sd, recvbuf, rc, rd);
  return MPI_SUCCESS;
}

int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm comm)
{
  const int n = comm->world->size;
  const int64_t b = (int64_t)count * type_size(t);
  std::vector<int64_t> sc(n, comm->rank == root ? b : 0), sd(n, 0), rc(n, 0), rd(n, 0);
  rc[root] = b;
  std::vector<char> tmp(b > 0 ? b : 1);
  a2av(comm, buf, sc, sd, tmp.data(), rc, rd);
  if (b > 0)
    std::memcpy(buf, tmp.data(), b);
  return MPI_SUCCESS;
}

// Contributions are folded in rank order 0..size-1 on every rank, so all ranks get
// the same bits and the order is a documented constant.
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype t, MPI_Op op, MPI_Comm comm)
{
  const int n = comm->world->size;
  const int64_t b = (int64_t)count * type_size(t);
  std::vector<char> all((size_t)std::max<int64_t>(b * n, 1));
  MPI_Allgather(sendbuf, count, t, all.data(), count, t, comm);
  if (b > 0)
    std::memcpy(recvbuf, all.data(), b);
  for (int r = 1; r < n; ++r)
    fold_any(recvbuf, all.data() + r * b, count, t, op);
  return MPI_SUCCESS;
}

int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype t, MPI_Op op, int root,
               MPI_Comm comm)
{
  std::vector<char> tmp((size_t)std::max<int64_t>((int64_t)count * type_size(t), 1));
  MPI_Allreduce(sendbuf, tmp.data(), count, t, op, comm);
  if (comm->rank == root)
    std::memcpy(recvbuf, tmp.data(), (size_t)count * type_size(t));
  return MPI_SUCCESS;
}

int MPI_Dist_graph_create_adjacent(MPI_Comm comm_old, int indegree, const int sources[], const int[],
                                   int outdegree, const int destinations[], const int[], MPI_Info, int,
                                   MPI_Comm* comm_dist_graph)
{
  auto* c = new spmv_shim_comm;
  c->world = comm_old->world;
  c->rank = comm_old->rank;
  c->is_graph = true;
  c->sources.assign(sources, sources + indegree);
  c->dests.assign(destinations, destinations + outdegree);
  *comm_dist_graph = c;
  return MPI_SUCCESS;
}

int MPI_Neighbor_alltoallv(const void* sendbuf, const int sendcounts[], const int sdispls[],
                           MPI_Datatype sendtype, void* recvbuf, const int recvcounts[], const int rdispls[],
                           MPI_Datatype recvtype, MPI_Comm comm)
{
  if (!comm->is_graph)
    die("MPI_Neighbor_alltoallv on a communicator without a graph topology");
  const int n = comm->world->size;
  const int64_t ss = type_size(sendtype), rs = type_size(recvtype);
  std::vector<int64_t> sc(n, 0), sd(n, 0), rc(n, 0), rd(n, 0);
  for (size_t i = 0; i < comm->dests.size(); ++i) {
    sc[comm->dests[i]] = sendcounts[i] * ss;
    sd[comm->dests[i]] = sdispls[i] * ss;
  }
  for (size_t j = 0; j < comm->sources.size(); ++j) {
    rc[comm->sources[j]] = recvcounts[j] * rs;
    rd[comm->sources[j]] = rdispls[j] * rs;
  }
  a2av(comm, sendbuf, sc, sd, recvbuf, rc, rd);
  return MPI_SUCCESS;
}

int MPI_Neighbor_alltoall(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf,
                          int recvcount, MPI_Datatype recvtype, MPI_Comm comm)
{
  std::vector<int> sc(comm->dests.size(), sendcount), sd(comm->dests.size());
  std::vector<int> rc(comm->sources.size(), recvcount), rd(comm->sources.size());
  for (size_t i = 0; i < sd.size(); ++i)
    sd[i] = (int)i * sendcount;
  for (size_t j = 0; j < rd.size(); ++j)
    rd[j] = (int)j * recvcount;
  return MPI_Neighbor_alltoallv(sendbuf, sc.data(), sd.data(), sendtype, recvbuf, rc.data(), rd.data(),
                                recvtype, comm);
}

// Completes inside the call (allowed: an MPI implementation may progress a
// non-blocking collective at any time); the request is a completed dummy.
int MPI_Ineighbor_alltoallv(const void* sendbuf, const int sendcounts[], const int sdispls[],
                            MPI_Datatype sendtype, void* recvbuf, const int recvcounts[], const int rdispls[],
                            MPI_Datatype recvtype, MPI_Comm comm, MPI_Request* request)
{
  MPI_Neighbor_alltoallv(sendbuf, sendcounts, sdispls, sendtype, recvbuf, recvcounts, rdispls, recvtype, comm);
  *request = MPI_REQUEST_NULL;
  return MPI_SUCCESS;
}

int MPI_Send(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm comm)
{
  comm->world->send(comm->rank, dest, tag, buf, (size_t)count * type_size(t));
  return MPI_SUCCESS;
}
int MPI_Recv(void* buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm comm, MPI_Status*)
{
  comm->world->recv(comm->rank, source, tag, buf, (size_t)count * type_size(t));
  return MPI_SUCCESS;
}
int MPI_Isend(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm comm,
              MPI_Request* request)
{
  MPI_Send(buf, count, t, dest, tag, comm); // eager
  *request = MPI_REQUEST_NULL;
  return MPI_SUCCESS;
}
int MPI_Irecv(void* buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm comm, MPI_Request* request)
{
  auto* r = new spmv_shim_request;
  r->comm = comm;
  r->buf = buf;
  r->bytes = (size_t)count * type_size(t);
  r->source = source;
  r->tag = tag;
  r->is_recv = true;
  *request = r;
  return MPI_SUCCESS;
}
int MPI_Wait(MPI_Request* request, MPI_Status*)
{
  if (request == nullptr || *request == MPI_REQUEST_NULL)
    return MPI_SUCCESS;
  spmv_shim_request* r = *request;
  if (r->is_recv)
    r->comm->world->recv(r->comm->rank, r->source, r->tag, r->buf, r->bytes);
  delete r;
  *request = MPI_REQUEST_NULL;
  return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request requests[], MPI_Status*)
{
  for (int i = 0; i < count; ++i)
    MPI_Wait(&requests[i], MPI_STATUS_IGNORE);
  return MPI_SUCCESS;
}

#define SHIM_UNSUPPORTED(name) die(name " is outside the supported subset (one-sided / shared-window models)")
int MPI_Comm_split_type(MPI_Comm, int, int, MPI_Info, MPI_Comm*) { SHIM_UNSUPPORTED("MPI_Comm_split_type"); }
int MPI_Comm_group(MPI_Comm, MPI_Group*) { SHIM_UNSUPPORTED("MPI_Comm_group"); }
int MPI_Group_translate_ranks(MPI_Group, int, const int[], MPI_Group, int[]) { SHIM_UNSUPPORTED("MPI_Group_translate_ranks"); }
int MPI_Info_create(MPI_Info* info)
{
  *info = 1;
  return MPI_SUCCESS;
}
int MPI_Info_set(MPI_Info, const char*, const char*) { return MPI_SUCCESS; }
int MPI_Win_create(void*, MPI_Aint, int, MPI_Info, MPI_Comm, MPI_Win*) { SHIM_UNSUPPORTED("MPI_Win_create"); }
int MPI_Win_allocate_shared(MPI_Aint, int, MPI_Info, MPI_Comm, void*, MPI_Win*) { SHIM_UNSUPPORTED("MPI_Win_allocate_shared"); }
int MPI_Win_shared_query(MPI_Win, int, MPI_Aint*, int*, void*) { SHIM_UNSUPPORTED("MPI_Win_shared_query"); }
int MPI_Win_free(MPI_Win*) { SHIM_UNSUPPORTED("MPI_Win_free"); }
int MPI_Win_fence(int, MPI_Win) { SHIM_UNSUPPORTED("MPI_Win_fence"); }
int MPI_Win_sync(MPI_Win) { SHIM_UNSUPPORTED("MPI_Win_sync"); }
int MPI_Win_lock(int, int, int, MPI_Win) { SHIM_UNSUPPORTED("MPI_Win_lock"); }
int MPI_Win_unlock(int, MPI_Win) { SHIM_UNSUPPORTED("MPI_Win_unlock"); }
int MPI_Win_lock_all(int, MPI_Win) { SHIM_UNSUPPORTED("MPI_Win_lock_all"); }
int MPI_Win_unlock_all(MPI_Win) { SHIM_UNSUPPORTED("MPI_Win_unlock_all"); }
int MPI_Put(const void*, int, MPI_Datatype, int, MPI_Aint, int, MPI_Datatype, MPI_Win) { SHIM_UNSUPPORTED("MPI_Put"); }

} // extern "C"
