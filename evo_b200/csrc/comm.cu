// NCCL behind the C-ABI (evo_comm_init / allreduce / destroy): the ONE collective of an EM iteration --
// the all-reduce of the packed sufficient statistics -- which replaces the reference's ~15 MPI Allreduce / Bcast
// calls per iteration (evo/models/bsc.py:230-231,257,269-274; evo/models/sssc.py:671-690; evo/utils/parallel.py:45-67).
// libnccl is not linked: it is resolved at run time (the copy already loaded by the host process, else
// libnccl.so.2 by name, else the path given to evo_comm_load), so the library also loads where NCCL is absent.
#include <dlfcn.h>

#include "common.cuh"

namespace evo {

typedef struct { char internal[128]; } nccl_uid;
typedef void *nccl_comm;
typedef int (*fn_get_uid)(nccl_uid *);
typedef int (*fn_init_rank)(nccl_comm *, int, nccl_uid, int);
typedef int (*fn_allreduce)(const void *, void *, size_t, int, int, nccl_comm, cudaStream_t);
typedef int (*fn_destroy)(nccl_comm);
typedef const char *(*fn_errstr)(int);

static void *g_nccl = nullptr;
static fn_get_uid p_get_uid;
static fn_init_rank p_init_rank;
static fn_allreduce p_allreduce;
static fn_destroy p_destroy;
static fn_errstr p_errstr;

static int nccl_load(const char *path) {
  if (g_nccl) return EVO_OK;
  void *h = nullptr;
  if (path && *path) h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    set_error("evo_comm: libnccl.so.2 not found (%s)", dlerror());
    return EVO_ERR_UNSUPPORTED;
  }
  p_get_uid = (fn_get_uid)dlsym(h, "ncclGetUniqueId");
  p_init_rank = (fn_init_rank)dlsym(h, "ncclCommInitRank");
  p_allreduce = (fn_allreduce)dlsym(h, "ncclAllReduce");
  p_destroy = (fn_destroy)dlsym(h, "ncclCommDestroy");
  p_errstr = (fn_errstr)dlsym(h, "ncclGetErrorString");
  if (!p_get_uid || !p_init_rank || !p_allreduce || !p_destroy) {
    set_error("evo_comm: NCCL symbols missing");
    return EVO_ERR_UNSUPPORTED;
  }
  g_nccl = h;
  return EVO_OK;
}
static int nccl_check(int rc, const char *what) {
  if (rc == 0) return EVO_OK;
  set_error("NCCL error %d (%s) in %s", rc, p_errstr ? p_errstr(rc) : "?", what);
  return EVO_ERR_CUDA;
}

}  // namespace evo

using namespace evo;

extern "C" int evo_comm_load(const char *libnccl_path) { return nccl_load(libnccl_path); }

extern "C" int evo_comm_unique_id(void *id128) {
  EVO_REQUIRE(id128 != nullptr, "null id");
  int rc = nccl_load(nullptr);
  if (rc) return rc;
  return nccl_check(p_get_uid(reinterpret_cast<nccl_uid *>(id128)), "ncclGetUniqueId");
}

extern "C" int evo_comm_init(int rank, int world, const void *id128, void **comm_out) {
  EVO_REQUIRE(id128 && comm_out && world >= 1 && rank >= 0 && rank < world, "bad arguments");
  int rc = nccl_load(nullptr);
  if (rc) return rc;
  nccl_uid id;
  memcpy(&id, id128, sizeof(id));
  nccl_comm c = nullptr;
  rc = nccl_check(p_init_rank(&c, world, id, rank), "ncclCommInitRank");
  if (rc) return rc;
  *comm_out = c;
  return EVO_OK;
}

extern "C" int evo_comm_allreduce(void *comm, double *buf, int64_t count, void *stream) {
  EVO_REQUIRE(comm && buf && count >= 0, "bad arguments");
  if (count == 0) return EVO_OK;
  // ncclFloat64 = 8, ncclSum = 0 (nccl.h); in place
  return nccl_check(p_allreduce(buf, buf, (size_t)count, 8, 0, comm, (cudaStream_t)stream), "ncclAllReduce");
}

extern "C" int evo_comm_destroy(void *comm) {
  if (!comm) return EVO_OK;
  EVO_REQUIRE(g_nccl != nullptr, "NCCL not loaded");
  return nccl_check(p_destroy(comm), "ncclCommDestroy");
}
