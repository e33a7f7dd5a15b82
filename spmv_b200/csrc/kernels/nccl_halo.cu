// nccl_halo.cu -- NCCL communicator, grouped send/recv halo exchange, scalar
// allreduce.
//
// Replaces the MPI message layer of the hot path: the Irecv/Send/Isend/Waitall
// loops of L2GMap::update_p2p / update_p2p_start (L2GMap.cpp:564-635), the
// neighbourhood alltoallv of update_collective (:502-526), and the 8-byte
// MPI_Allreduce calls of cg.cpp:49,65,75.  One rank per GPU; every call is
// enqueued on a CUDA stream, nothing blocks the host.
//
// NCCL is bound at run time (dlopen of libnccl.so.2): in a PyTorch process that
// is the copy torch already loaded, otherwise the system one.  Only the stable
// core API is used, so 2.27 headers against a 2.28 library is fine.
#include "common.cuh"

#include <dlfcn.h>
#include <mutex>
#include <nccl.h>
#include <string.h>

namespace b200
{

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t)
      = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

static NcclApi& nccl()
{
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle)
        break;
    }
    if (!api.handle)
      return;
#define B200_SYM(field, name)                                                  \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name));  \
  if (!api.field)                                                              \
    return;
    B200_SYM(GetUniqueId, "ncclGetUniqueId")
    B200_SYM(CommInitRank, "ncclCommInitRank")
    B200_SYM(CommDestroy, "ncclCommDestroy")
    B200_SYM(GroupStart, "ncclGroupStart")
    B200_SYM(GroupEnd, "ncclGroupEnd")
    B200_SYM(Send, "ncclSend")
    B200_SYM(Recv, "ncclRecv")
    B200_SYM(AllReduce, "ncclAllReduce")
    B200_SYM(GetErrorString, "ncclGetErrorString")
#undef B200_SYM
    api.ok = true;
  });
  return api;
}

static int nccl_fail(ncclResult_t r, const char* what)
{
  NcclApi& api = nccl();
  set_error("NCCL error %d (%s) in %s", (int)r,
            api.GetErrorString ? api.GetErrorString(r) : "?", what);
  return B200_ENCCL;
}

#define B200_NCCL(call)                                                        \
  do {                                                                         \
    ncclResult_t r__ = (call);                                                 \
    if (r__ != ncclSuccess)                                                    \
      return nccl_fail(r__, #call);                                            \
  } while (0)

static int need_nccl()
{
  if (!nccl().ok) {
    set_error("libnccl.so.2 could not be loaded: %s", dlerror());
    return B200_ENCCL;
  }
  return B200_OK;
}

} // namespace b200

using namespace b200;

static_assert(sizeof(ncclUniqueId) == B200_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");

extern "C" {

int b200_nccl_unique_id(char id[B200_NCCL_UNIQUE_ID_BYTES])
{
  B200_REQUIRE(id != nullptr, "id is NULL");
  if (int rc = need_nccl())
    return rc;
  ncclUniqueId uid;
  B200_NCCL(nccl().GetUniqueId(&uid));
  memcpy(id, &uid, sizeof(uid));
  return B200_OK;
}

int b200_nccl_comm_create(void** comm, int nranks, int rank,
                          const char id[B200_NCCL_UNIQUE_ID_BYTES])
{
  B200_REQUIRE(comm != nullptr && id != nullptr, "NULL argument");
  B200_REQUIRE(nranks > 0 && rank >= 0 && rank < nranks, "bad rank / nranks");
  if (int rc = need_nccl())
    return rc;
  ncclUniqueId uid;
  memcpy(&uid, id, sizeof(uid));
  ncclComm_t c = nullptr;
  B200_NCCL(nccl().CommInitRank(&c, nranks, uid, rank));
  *comm = c;
  return B200_OK;
}

int b200_nccl_comm_destroy(void* comm)
{
  if (comm == nullptr)
    return B200_OK;
  if (int rc = need_nccl())
    return rc;
  B200_NCCL(nccl().CommDestroy(static_cast<ncclComm_t>(comm)));
  return B200_OK;
}

int b200_halo_exchange(void* comm, int nnbr, const int32_t* nbr_rank, const void* d_sendbuf,
                       const int32_t* send_cnt, const int32_t* send_off, void* d_recvbuf,
                       const int32_t* recv_cnt, const int32_t* recv_off, int elem_bytes,
                       void* comm_stream)
{
  B200_REQUIRE(nnbr >= 0, "negative neighbour count");
  if (nnbr == 0)
    return B200_OK;
  B200_REQUIRE(comm != nullptr, "comm is NULL");
  B200_REQUIRE(nbr_rank && send_cnt && send_off && recv_cnt && recv_off, "NULL plan array");
  B200_REQUIRE(elem_bytes == 4 || elem_bytes == 8, "elem_bytes must be 4 or 8");
  if (int rc = need_nccl())
    return rc;
  NcclApi& api = nccl();
  ncclComm_t c = static_cast<ncclComm_t>(comm);
  cudaStream_t st = as_stream(comm_stream);
  const ncclDataType_t dt = elem_bytes == 8 ? ncclDouble : ncclFloat;
  const char* sb = static_cast<const char*>(d_sendbuf);
  char* rb = static_cast<char*>(d_recvbuf);
  B200_NCCL(api.GroupStart());
  // an error inside the group must still close it: NCCL keeps the group open per thread otherwise
  ncclResult_t bad = ncclSuccess;
  for (int i = 0; i < nnbr && bad == ncclSuccess; ++i) {
    if (recv_cnt[i] > 0)
      bad = api.Recv(rb + (size_t)recv_off[i] * elem_bytes, (size_t)recv_cnt[i], dt, nbr_rank[i], c, st);
    if (bad == ncclSuccess && send_cnt[i] > 0)
      bad = api.Send(sb + (size_t)send_off[i] * elem_bytes, (size_t)send_cnt[i], dt, nbr_rank[i], c, st);
  }
  const ncclResult_t end = api.GroupEnd();
  B200_NCCL(bad);
  B200_NCCL(end);
  return B200_OK;
}

int b200_allreduce_sum_f64(void* comm, double* d_inout, int32_t n, void* stream)
{
  B200_REQUIRE(n >= 0, "negative count");
  if (n == 0 || comm == nullptr) // single rank: the local value is the sum
    return B200_OK;
  if (int rc = need_nccl())
    return rc;
  B200_NCCL(nccl().AllReduce(d_inout, d_inout, (size_t)n, ncclDouble, ncclSum,
                             static_cast<ncclComm_t>(comm), as_stream(stream)));
  return B200_OK;
}

} // extern "C"
