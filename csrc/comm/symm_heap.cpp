// Native symmetric heap with an NVLS multicast alias ("memory model").
//
// Every rank creates ONE physical allocation with the CUDA virtual-memory-management API
// (cuMemCreate, shareable as a POSIX file descriptor), the descriptors travel between the rank
// processes over Unix sockets (SCM_RIGHTS: the MPI shim's socketpairs for bin/jordan, abstract Unix
// sockets under torchrun), and every rank maps every peer's allocation into its own address space
// (cuMemImportFromShareableHandle / cuMemMap / cuMemSetAccess).  On NVSwitch systems rank 0
// additionally creates a multicast object (cuMulticastCreate), every rank adds its device and binds
// its allocation (cuMulticastAddDevice / cuMulticastBindMem) and maps the object: a store to that
// alias (multimem.st) is replicated by the switch into every GPU's copy.  This replaces the
// reference's per-step MPI_Bcast of the pivot row (main.cpp:1097) and the mailbox MPI_Allreduce
// (main.cpp:1074) with plain stores issued from inside the kernels.
//
// The API is staged so that the CALLER moves the descriptors with whatever channel it has:
//
//   jb_symm_heap_create      local allocation, returns the descriptor to hand to the peers
//   jb_symm_heap_import      descriptors of all ranks (own slot ignored) -> peer mappings
//   jb_symm_heap_mc_create   rank 0: multicast object, returns its descriptor (-1: no multicast)
//   jb_symm_heap_mc_import   ranks != 0: descriptor received from rank 0
//   jb_symm_heap_mc_add      every rank;   -- barrier --
//   jb_symm_heap_mc_bind     every rank;   -- barrier --   (maps the alias as well)
//
// All driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the library still
// loads on a machine without libcuda (the CPU-only build box).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <vector>

#include "../../include/jordan_b200.h"

void jb_set_error(const char* fmt, ...);

namespace {

struct Drv {
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t,
                               size_t, unsigned long long);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok = false;
};

template <typename F>
bool load_one(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p) {
    jb_set_error("symm_heap: driver entry point %s not available", name);
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

Drv* drv() {
  static Drv d;
  if (d.ok) return &d;
  bool ok = load_one(d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") &&
            load_one(d.MemCreate, "cuMemCreate") && load_one(d.MemRelease, "cuMemRelease") &&
            load_one(d.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
            load_one(d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
            load_one(d.MemAddressReserve, "cuMemAddressReserve") &&
            load_one(d.MemAddressFree, "cuMemAddressFree") && load_one(d.MemMap, "cuMemMap") &&
            load_one(d.MemUnmap, "cuMemUnmap") && load_one(d.MemSetAccess, "cuMemSetAccess") &&
            load_one(d.MulticastCreate, "cuMulticastCreate") &&
            load_one(d.MulticastAddDevice, "cuMulticastAddDevice") &&
            load_one(d.MulticastBindMem, "cuMulticastBindMem") &&
            load_one(d.MulticastUnbind, "cuMulticastUnbind") &&
            load_one(d.MulticastGetGranularity, "cuMulticastGetGranularity") &&
            load_one(d.DeviceGet, "cuDeviceGet") && load_one(d.DeviceGetAttribute, "cuDeviceGetAttribute") &&
            load_one(d.GetErrorString, "cuGetErrorString");
  d.ok = ok;
  return ok ? &d : nullptr;
}

const char* errstr(Drv* d, CUresult r) {
  const char* s = nullptr;
  if (d && d->GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUDA driver error";
}

#define DRV_CHECK(d, call)                                                             \
  do {                                                                                 \
    CUresult _r = (call);                                                              \
    if (_r != CUDA_SUCCESS) {                                                          \
      jb_set_error("symm_heap: %s failed: %s (%d)", #call, errstr(d, _r), (int)_r);    \
      return JB_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

}  // namespace

struct JbSymmHeap {
  int rank = 0, size = 1, device = 0;
  CUdevice cudev = 0;
  size_t bytes = 0;       // mapped size (request rounded up to the granularity)
  size_t requested = 0;
  size_t gran = 0;
  CUmemGenericAllocationHandle local = 0;
  std::vector<CUmemGenericAllocationHandle> peer_handles;  // imported (own slot unused)
  std::vector<CUdeviceptr> va;                             // va[r]: rank r's memory as seen from here
  bool mc_supported = false;
  CUmemGenericAllocationHandle mc = 0;
  bool mc_added = false, mc_bound = false;
  CUdeviceptr mc_va = 0;
};

namespace {

int map_rw(Drv* d, JbSymmHeap* h, CUdeviceptr* out, CUmemGenericAllocationHandle handle) {
  CUdeviceptr p = 0;
  DRV_CHECK(d, d->MemAddressReserve(&p, h->bytes, h->gran, 0, 0));
  CUresult r = d->MemMap(p, h->bytes, 0, handle, 0);
  if (r != CUDA_SUCCESS) {
    d->MemAddressFree(p, h->bytes);
    jb_set_error("symm_heap: cuMemMap failed: %s (%d)", errstr(d, r), (int)r);
    return JB_ERR_CUDA;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = h->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d->MemSetAccess(p, h->bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d->MemUnmap(p, h->bytes);
    d->MemAddressFree(p, h->bytes);
    jb_set_error("symm_heap: cuMemSetAccess failed: %s (%d) - no peer access between the GPUs?",
                 errstr(d, r), (int)r);
    return JB_ERR_CUDA;
  }
  *out = p;
  return JB_OK;
}

}  // namespace

extern "C" {

int jb_symm_heap_create(size_t bytes, int device, int rank, int size, int want_multicast,
                        JbSymmHeap** out, int* export_fd) {
  if (!out || !export_fd || bytes == 0 || size < 1 || rank < 0 || rank >= size) {
    jb_set_error("jb_symm_heap_create: bad arguments");
    return JB_ERR_ARG;
  }
  *out = nullptr;
  *export_fd = -1;
  if (cudaSetDevice(device) != cudaSuccess || cudaFree(0) != cudaSuccess) {  // primary context current
    jb_set_error("jb_symm_heap_create: cannot select device %d", device);
    return JB_ERR_CUDA;
  }
  Drv* d = drv();
  if (!d) return JB_ERR_CUDA;
  JbSymmHeap* h = new JbSymmHeap();
  h->rank = rank; h->size = size; h->device = device; h->requested = bytes;
  h->peer_handles.assign(size, 0);
  h->va.assign(size, 0);
  int rc = JB_OK;
  do {
    CUresult r;
    if ((r = d->DeviceGet(&h->cudev, device)) != CUDA_SUCCESS) { rc = JB_ERR_CUDA; break; }
    int vmm = 0, fdok = 0, mcs = 0;
    d->DeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, h->cudev);
    d->DeviceGetAttribute(&fdok, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, h->cudev);
    d->DeviceGetAttribute(&mcs, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, h->cudev);
    if (!vmm || !fdok) {
      jb_set_error("symm_heap: device %d lacks VMM / POSIX-fd shareable handles", device);
      rc = JB_ERR_CUDA;
      break;
    }
    h->mc_supported = want_multicast && mcs && size > 1;
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran = 0;
    if ((r = d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED)) != CUDA_SUCCESS) {
      jb_set_error("symm_heap: cuMemGetAllocationGranularity: %s", errstr(d, r));
      rc = JB_ERR_CUDA;
      break;
    }
    if (h->mc_supported) {  // a multicast binding wants its own (possibly larger) granularity
      CUmulticastObjectProp mp;
      memset(&mp, 0, sizeof(mp));
      mp.numDevices = (unsigned)size;
      mp.size = bytes;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
        gran = mg;
    }
    h->gran = gran;
    h->bytes = (bytes + gran - 1) / gran * gran;
    if ((r = d->MemCreate(&h->local, h->bytes, &prop, 0)) != CUDA_SUCCESS) {
      jb_set_error("Not enough memory! (symmetric heap, %zu bytes: %s)", h->bytes, errstr(d, r));
      rc = JB_ERR_NOMEM;
      break;
    }
    int fd = -1;
    if ((r = d->MemExportToShareableHandle(&fd, h->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0)) !=
        CUDA_SUCCESS) {
      jb_set_error("symm_heap: cuMemExportToShareableHandle: %s", errstr(d, r));
      rc = JB_ERR_CUDA;
      break;
    }
    if ((rc = map_rw(d, h, &h->va[rank], h->local)) != JB_OK) {
      close(fd);
      break;
    }
    if (cudaMemset((void*)h->va[rank], 0, h->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
      jb_set_error("symm_heap: clearing the heap failed");
      close(fd);
      rc = JB_ERR_CUDA;
      break;
    }
    *export_fd = fd;
  } while (0);
  if (rc != JB_OK) {
    jb_symm_heap_destroy(h);
    return rc;
  }
  *out = h;
  return JB_OK;
}

// fds[r] = descriptor exported by rank r (fds[rank] is ignored).  The descriptors are closed here.
int jb_symm_heap_import(JbSymmHeap* h, const int* fds) {
  Drv* d = drv();
  if (!d || !h) return JB_ERR_ARG;
  cudaSetDevice(h->device);
  int rc = JB_OK;
  for (int r = 0; r < h->size; ++r) {
    if (r == h->rank) continue;
    if (fds[r] < 0) {
      if (rc == JB_OK) {
        jb_set_error("symm_heap: no descriptor for rank %d", r);
        rc = JB_ERR_ARG;
      }
      continue;
    }
    if (rc != JB_OK) {  // an earlier peer failed: still close what we were handed
      close(fds[r]);
      continue;
    }
    CUresult cr = d->MemImportFromShareableHandle(&h->peer_handles[r], (void*)(uintptr_t)fds[r],
                                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    close(fds[r]);
    if (cr != CUDA_SUCCESS) {
      jb_set_error("symm_heap: import of rank %d's allocation failed: %s", r, errstr(d, cr));
      rc = JB_ERR_CUDA;
      continue;
    }
    rc = map_rw(d, h, &h->va[r], h->peer_handles[r]);
  }
  return rc;
}

int jb_symm_heap_multicast_supported(const JbSymmHeap* h) { return h && h->mc_supported ? 1 : 0; }

// rank 0 only.  *export_fd = descriptor of the multicast object for the other ranks.
int jb_symm_heap_mc_create(JbSymmHeap* h, int* export_fd) {
  Drv* d = drv();
  *export_fd = -1;
  if (!d || !h || !h->mc_supported) return JB_ERR_ARG;
  cudaSetDevice(h->device);
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)h->size;
  mp.size = h->bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DRV_CHECK(d, d->MulticastCreate(&h->mc, &mp));
  int fd = -1;
  DRV_CHECK(d, d->MemExportToShareableHandle(&fd, h->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *export_fd = fd;
  return JB_OK;
}

int jb_symm_heap_mc_import(JbSymmHeap* h, int fd) {
  Drv* d = drv();
  if (!d || !h || fd < 0) return JB_ERR_ARG;
  cudaSetDevice(h->device);
  CUresult cr = d->MemImportFromShareableHandle(&h->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);
  if (cr != CUDA_SUCCESS) {
    jb_set_error("symm_heap: import of the multicast object failed: %s", errstr(d, cr));
    return JB_ERR_CUDA;
  }
  return JB_OK;
}

int jb_symm_heap_mc_add(JbSymmHeap* h) {
  Drv* d = drv();
  if (!d || !h || !h->mc) return JB_ERR_ARG;
  cudaSetDevice(h->device);
  DRV_CHECK(d, d->MulticastAddDevice(h->mc, h->cudev));
  h->mc_added = true;
  return JB_OK;
}

// after EVERY rank has added its device (barrier): bind the local allocation and map the alias
int jb_symm_heap_mc_bind(JbSymmHeap* h) {
  Drv* d = drv();
  if (!d || !h || !h->mc_added) return JB_ERR_ARG;
  cudaSetDevice(h->device);
  DRV_CHECK(d, d->MulticastBindMem(h->mc, 0, h->local, 0, h->bytes, 0));
  h->mc_bound = true;
  return map_rw(d, h, &h->mc_va, h->mc);
}

// give up the multicast alias (some rank could not set it up): stores fall back to unicast
void jb_symm_heap_mc_drop(JbSymmHeap* h) {
  Drv* d = drv();
  if (!d || !h) return;
  if (h->mc_va) {
    d->MemUnmap(h->mc_va, h->bytes);
    d->MemAddressFree(h->mc_va, h->bytes);
    h->mc_va = 0;
  }
  if (h->mc_bound) d->MulticastUnbind(h->mc, h->cudev, 0, h->bytes);
  h->mc_bound = false;
  if (h->mc) d->MemRelease(h->mc);
  h->mc = 0;
  h->mc_added = false;
  h->mc_supported = false;
}

void* jb_symm_heap_ptr(const JbSymmHeap* h, int r) {
  return (h && r >= 0 && r < h->size) ? (void*)h->va[r] : nullptr;
}
void* jb_symm_heap_mc_ptr(const JbSymmHeap* h) { return h ? (void*)h->mc_va : nullptr; }
size_t jb_symm_heap_bytes(const JbSymmHeap* h) { return h ? h->bytes : 0; }

void jb_symm_heap_destroy(JbSymmHeap* h) {
  if (!h) return;
  Drv* d = drv();
  if (d) {
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    jb_symm_heap_mc_drop(h);
    for (int r = 0; r < h->size; ++r) {
      if (h->va[r]) {
        d->MemUnmap(h->va[r], h->bytes);
        d->MemAddressFree(h->va[r], h->bytes);
      }
      if (r != h->rank && h->peer_handles[r]) d->MemRelease(h->peer_handles[r]);
    }
    if (h->local) d->MemRelease(h->local);
  }
  delete h;
}

}  // extern "C"
