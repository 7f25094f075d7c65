// Mesh-axis communicators and collectives of the C ABI: what the case scripts get
// from `Mesh(create_device_mesh((2,4)), ('data','model'))` plus the psum / all_gather / psum_scatter / all_to_all XLA
// inserts around a sharded dot (case1a.py:15,49; case6_attention.py:155-156,229-237), for a binder that runs one process
// per GPU and has no torch.distributed.  Thin, on purpose: NCCL over NVLink 5 / NVSwitch does the transport; this file
// owns the mesh arithmetic (row-major device ids, an axis communicator = the ranks that differ only in that coordinate,
// ordered by it) and the error / stream conventions of the rest of the ABI.  The fused GEMM -> reduce-scatter of the hot
// layers does NOT go through here (csrc/peer.cu: peer-memory pushes from the GEMM epilogue).
//
// NCCL is resolved at run time (dlopen of libnccl.so.2: the copy already in the process if the host framework loaded
// one), so the library has no link-time dependency on it and still loads on a box without NCCL; every entry point then
// fails with B200_ERR_UNSUPPORTED.
#include <dlfcn.h>
#include <stddef.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "host_common.h"

namespace b200 {

// The slice of nccl.h this file needs (NCCL's C ABI is stable across 2.x for these entry points).
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { kNcclSum = 0, kNcclFloat32 = 7, kNcclBfloat16 = 9 };

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t*, void*);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*ReduceScatter)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  const char* (*GetErrorString)(ncclResult_t);
  bool ok = false;
};

static NcclApi& nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
    bool all = true;
    auto sym = [&](const char* name) {
      void* p = dlsym(h, name);
      if (!p) all = false;
      return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommSplit = reinterpret_cast<decltype(api.CommSplit)>(sym("ncclCommSplit"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.ReduceScatter = reinterpret_cast<decltype(api.ReduceScatter)>(sym("ncclReduceScatter"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    api.ok = all;
  });
  return api;
}

#define B200_NCCL_READY()                                                                              \
  do {                                                                                                 \
    if (!nccl().ok) return fail(B200_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded in this process"); \
  } while (0)

#define B200_NCCL_CHECK(expr)                                                                        \
  do {                                                                                               \
    ncclResult_t _r = (expr);                                                                        \
    if (_r != 0)                                                                                     \
      return fail(B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, nccl().GetErrorString(_r), __FILE__, __LINE__); \
  } while (0)

}  // namespace b200

using namespace b200;

// One rank's view of the mesh: its coordinates and, per axis, the communicator of the ranks that share all its other
// coordinates (created on first use with ncclCommSplit; color = the index of that line of the mesh, key = the coordinate).
struct b200_mesh {
  int world = 0, rank = 0, naxes = 0;
  int shape[8] = {0};
  int coord[8] = {0};
  ncclComm_t world_comm = nullptr;
  ncclComm_t axis_comm[8] = {nullptr};
};
struct b200_comm {
  ncclComm_t comm;
  int size, rank;
};

static int dtype_of(int dtype, int* nccl_type, size_t* elem) {
  if (dtype == B200_BF16) { *nccl_type = kNcclBfloat16; *elem = 2; return B200_OK; }
  if (dtype == B200_F32) { *nccl_type = kNcclFloat32; *elem = 4; return B200_OK; }
  return fail(B200_ERR_INVALID_ARGUMENT, "collective dtype must be B200_BF16 or B200_F32 (got %d)", dtype);
}

extern "C" int b200_comm_unique_id(void* out128) {
  B200_REQUIRE(out128, "b200_comm_unique_id: null output");
  B200_NCCL_READY();
  ncclUniqueId id;
  B200_NCCL_CHECK(nccl().GetUniqueId(&id));
  memcpy(out128, id.internal, sizeof(id.internal));
  return B200_OK;
}

extern "C" int b200_mesh_create(const void* unique_id128, int world, int rank, const int* shape, int naxes,
                                b200_mesh** out) {
  B200_REQUIRE(unique_id128 && shape && out, "b200_mesh_create: null argument");
  B200_REQUIRE(naxes >= 1 && naxes <= 8, "b200_mesh_create: 1..8 mesh axes (got %d)", naxes);
  int n = 1;
  for (int i = 0; i < naxes; ++i) {
    B200_REQUIRE(shape[i] >= 1, "b200_mesh_create: axis %d has size %d", i, shape[i]);
    n *= shape[i];
  }
  B200_REQUIRE(n == world, "b200_mesh_create: mesh of %d devices for a job of %d ranks", n, world);
  B200_REQUIRE(rank >= 0 && rank < world, "b200_mesh_create: rank %d out of range", rank);
  B200_NCCL_READY();
  b200_mesh* m = new b200_mesh();
  m->world = world;
  m->rank = rank;
  m->naxes = naxes;
  int r = rank;                                  // row-major: the last axis varies fastest (device id = 4x + y on 2x4)
  for (int i = naxes - 1; i >= 0; --i) {
    m->shape[i] = shape[i];
    m->coord[i] = r % shape[i];
    r /= shape[i];
  }
  ncclUniqueId id;
  memcpy(id.internal, unique_id128, sizeof(id.internal));
  ncclResult_t rc = nccl().CommInitRank(&m->world_comm, world, id, rank);
  if (rc != 0) {
    delete m;
    return fail(B200_ERR_CUDA, "ncclCommInitRank failed: %s", nccl().GetErrorString(rc));
  }
  *out = m;
  return B200_OK;
}

extern "C" int b200_mesh_axis_comm(b200_mesh* mesh, int axis, b200_comm** out) {
  B200_REQUIRE(mesh && out, "b200_mesh_axis_comm: null argument");
  B200_REQUIRE(axis >= 0 && axis < mesh->naxes, "b200_mesh_axis_comm: axis %d of a %d-axis mesh", axis, mesh->naxes);
  B200_NCCL_READY();
  if (!mesh->axis_comm[axis]) {
    int color = 0;                               // index of this rank's line along `axis`: its other coordinates
    for (int i = 0; i < mesh->naxes; ++i)
      if (i != axis) color = color * mesh->shape[i] + mesh->coord[i];
    B200_NCCL_CHECK(nccl().CommSplit(mesh->world_comm, color, mesh->coord[axis], &mesh->axis_comm[axis], nullptr));
  }
  b200_comm* c = new b200_comm();
  c->comm = mesh->axis_comm[axis];
  c->size = mesh->shape[axis];
  c->rank = mesh->coord[axis];
  *out = c;
  return B200_OK;
}

extern "C" int b200_comm_size(const b200_comm* comm) { return comm ? comm->size : -1; }
extern "C" int b200_comm_rank(const b200_comm* comm) { return comm ? comm->rank : -1; }
extern "C" int b200_comm_release(b200_comm* comm) {   // the NCCL communicator itself belongs to the mesh
  delete comm;
  return B200_OK;
}

extern "C" int b200_mesh_destroy(b200_mesh* mesh) {
  if (!mesh) return B200_OK;
  B200_NCCL_READY();
  for (int i = 0; i < mesh->naxes; ++i)
    if (mesh->axis_comm[i]) nccl().CommDestroy(mesh->axis_comm[i]);
  if (mesh->world_comm) nccl().CommDestroy(mesh->world_comm);
  delete mesh;
  return B200_OK;
}

// recv[i] = sum over the axis of send[i]; in place allowed (psum over a mesh axis).
extern "C" int b200_all_reduce(b200_comm* comm, const void* send, void* recv, int64_t count, int dtype, void* stream) {
  B200_REQUIRE(comm && send && recv && count >= 0, "b200_all_reduce: bad arguments");
  int t; size_t e;
  if (int rc = dtype_of(dtype, &t, &e)) return rc;
  B200_NCCL_READY();
  B200_NCCL_CHECK(nccl().AllReduce(send, recv, static_cast<size_t>(count), t, kNcclSum, comm->comm,
                                   reinterpret_cast<cudaStream_t>(stream)));
  return B200_OK;
}

// send holds size() consecutive blocks of recv_count elements; rank r receives the sum of everybody's block r
// (psum_scatter: the reduction a row-parallel layer needs when its output stays sharded).
extern "C" int b200_reduce_scatter(b200_comm* comm, const void* send, void* recv, int64_t recv_count, int dtype,
                                   void* stream) {
  B200_REQUIRE(comm && send && recv && recv_count >= 0, "b200_reduce_scatter: bad arguments");
  int t; size_t e;
  if (int rc = dtype_of(dtype, &t, &e)) return rc;
  B200_NCCL_READY();
  B200_NCCL_CHECK(nccl().ReduceScatter(send, recv, static_cast<size_t>(recv_count), t, kNcclSum, comm->comm,
                                       reinterpret_cast<cudaStream_t>(stream)));
  return B200_OK;
}

// recv = the ranks' send buffers of send_count elements, concatenated in axis order (FSDP weight gather, K/V gather).
extern "C" int b200_all_gather(b200_comm* comm, const void* send, void* recv, int64_t send_count, int dtype, void* stream) {
  B200_REQUIRE(comm && send && recv && send_count >= 0, "b200_all_gather: bad arguments");
  int t; size_t e;
  if (int rc = dtype_of(dtype, &t, &e)) return rc;
  B200_NCCL_READY();
  B200_NCCL_CHECK(nccl().AllGather(send, recv, static_cast<size_t>(send_count), t, comm->comm,
                                   reinterpret_cast<cudaStream_t>(stream)));
  return B200_OK;
}

// Block p of send (count elements each) goes to rank p, which stores the block from rank q at block q of recv
// (the M-sharded -> S-sharded re-split after the literal out-projection, case6_attention.py:141).
extern "C" int b200_all_to_all(b200_comm* comm, const void* send, void* recv, int64_t count, int dtype, void* stream) {
  B200_REQUIRE(comm && send && recv && count >= 0, "b200_all_to_all: bad arguments");
  B200_REQUIRE(send != recv, "b200_all_to_all: in-place exchange is not supported");
  int t; size_t e;
  if (int rc = dtype_of(dtype, &t, &e)) return rc;
  B200_NCCL_READY();
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  B200_NCCL_CHECK(nccl().GroupStart());
  for (int p = 0; p < comm->size; ++p) {
    const char* s = static_cast<const char*>(send) + static_cast<size_t>(p) * count * e;
    char* r = static_cast<char*>(recv) + static_cast<size_t>(p) * count * e;
    ncclResult_t rc = nccl().Send(s, static_cast<size_t>(count), t, p, comm->comm, st);
    if (rc == 0) rc = nccl().Recv(r, static_cast<size_t>(count), t, p, comm->comm, st);
    if (rc != 0) {
      nccl().GroupEnd();
      return fail(B200_ERR_CUDA, "ncclSend/ncclRecv failed: %s", nccl().GetErrorString(rc));
    }
  }
  B200_NCCL_CHECK(nccl().GroupEnd());
  return B200_OK;
}
