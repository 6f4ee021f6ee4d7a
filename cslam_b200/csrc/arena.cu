#include "arena.h"

#include <stdlib.h>
#include <string.h>

namespace cslam {

namespace {
template <typename F>
bool load_sym(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || fn == nullptr) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}
size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
}  // namespace

const DriverApi& driver_api() {
  static DriverApi api = [] {
    DriverApi a;
    bool ok = true;
    ok &= load_sym("cuMemAddressReserve", &a.memAddressReserve);
    ok &= load_sym("cuMemAddressFree", &a.memAddressFree);
    ok &= load_sym("cuMemCreate", &a.memCreate);
    ok &= load_sym("cuMemRelease", &a.memRelease);
    ok &= load_sym("cuMemMap", &a.memMap);
    ok &= load_sym("cuMemUnmap", &a.memUnmap);
    ok &= load_sym("cuMemSetAccess", &a.memSetAccess);
    ok &= load_sym("cuMemGetAllocationGranularity", &a.memGetAllocationGranularity);
    a.ok = ok;
    load_sym("cuTensorMapEncodeTiled", &a.tensorMapEncodeTiled);
    return a;
  }();
  return api;
}

Arena::~Arena() { release(); }

void Arena::release() {
  if (!base_) return;
  if (vmm_) {
    const DriverApi& d = driver_api();
    size_t off = 0;
    for (auto& c : chunks_) {
      d.memUnmap((CUdeviceptr)base_ + off, c.size);
      d.memRelease(c.handle);
      off += c.size;
    }
    d.memAddressFree((CUdeviceptr)base_, reserved_);
    chunks_.clear();
  } else {
    cudaFree(base_);
  }
  base_ = nullptr;
  mapped_ = reserved_ = 0;
}

bool Arena::ensure(int device, size_t bytes, cudaStream_t stream, std::string* err) {
  if (bytes <= mapped_) return true;
  if (!decided_) {
    const char* no = getenv("CSLAM_NNS_NO_VMM");
    vmm_ = driver_api().ok && !(no && no[0] == '1');
    if (vmm_) {
      CUmemAllocationProp prop;
      memset(&prop, 0, sizeof(prop));
      prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
      prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      prop.location.id = device;
      if (driver_api().memGetAllocationGranularity(&gran_, &prop,
                                                   CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS ||
          gran_ == 0)
        vmm_ = false;
    }
    decided_ = true;
  }
  return vmm_ ? ensure_vmm(device, bytes, stream, err) : ensure_malloc(bytes, stream, err);
}

bool Arena::ensure_vmm(int device, size_t bytes, cudaStream_t stream, std::string* err) {
  const DriverApi& d = driver_api();
  const size_t want = round_up(bytes, gran_);

  // 1. virtual range: reserve generously (4x the request, >= 256 MB); when it runs out, move the
  //    existing physical chunks to a larger range — a re-map, not a copy.
  if (want > reserved_) {
    size_t new_res = round_up(want * 4 > (size_t(256) << 20) ? want * 4 : (size_t(256) << 20), gran_);
    CUdeviceptr nb = 0;
    CUresult r = d.memAddressReserve(&nb, new_res, 0, 0, 0);
    if (r != CUDA_SUCCESS) {
      new_res = want;
      r = d.memAddressReserve(&nb, new_res, 0, 0, 0);
    }
    if (r != CUDA_SUCCESS) {
      *err = "cuMemAddressReserve failed (" + std::to_string((int)r) + ")";
      return false;
    }
    if (base_) {
      // Transactional: the chunks are mapped into the NEW range first (a physical allocation may be mapped
      // at two addresses), and only when all of them are in place is the old range torn down.  A failure
      // rolls the new range back and leaves the arena exactly as it was.
      size_t off = 0;
      bool ok = true;
      for (auto& c : chunks_) {
        if (d.memMap(nb + off, c.size, 0, c.handle, 0) != CUDA_SUCCESS) {
          ok = false;
          break;
        }
        off += c.size;
      }
      CUmemAccessDesc acc;
      memset(&acc, 0, sizeof(acc));
      acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      acc.location.id = device;
      acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      if (ok && off && d.memSetAccess(nb, off, &acc, 1) != CUDA_SUCCESS) ok = false;
      if (!ok) {
        size_t undo = 0;
        for (auto& c : chunks_) {
          if (undo >= off) break;
          d.memUnmap(nb + undo, c.size);
          undo += c.size;
        }
        d.memAddressFree(nb, new_res);
        *err = "cuMemMap / cuMemSetAccess (re-map into a larger range) failed";
        return false;
      }
      cudaDeviceSynchronize();  // nothing may be running against the old range when it goes away
      size_t old_off = 0;
      for (auto& c : chunks_) {
        d.memUnmap((CUdeviceptr)base_ + old_off, c.size);
        old_off += c.size;
      }
      d.memAddressFree((CUdeviceptr)base_, reserved_);
    }
    base_ = (void*)nb;
    reserved_ = new_res;
  }

  // 2. one new physical chunk for the growth
  const size_t add = want - mapped_;
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  Chunk c;
  c.size = add;
  CUresult r = d.memCreate(&c.handle, add, &prop, 0);
  if (r != CUDA_SUCCESS) {
    *err = "cuMemCreate(" + std::to_string(add) + " B) failed (" + std::to_string((int)r) + ")";
    return false;
  }
  r = d.memMap((CUdeviceptr)base_ + mapped_, add, 0, c.handle, 0);
  if (r != CUDA_SUCCESS) {
    d.memRelease(c.handle);
    *err = "cuMemMap failed (" + std::to_string((int)r) + ")";
    return false;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.memSetAccess((CUdeviceptr)base_ + mapped_, add, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.memUnmap((CUdeviceptr)base_ + mapped_, add);
    d.memRelease(c.handle);
    *err = "cuMemSetAccess failed (" + std::to_string((int)r) + ")";
    return false;
  }
  chunks_.push_back(c);
  // rows past n read as zeros in the reference (torch.zeros, nns_matching.py:36,38) and the pad
  // lanes of every row must be zero for the kernels
  cudaError_t e = cudaMemsetAsync((char*)base_ + mapped_, 0, add, stream);
  if (e != cudaSuccess) {
    *err = std::string("cudaMemsetAsync: ") + cudaGetErrorString(e);
    return false;
  }
  mapped_ = want;
  return true;
}

bool Arena::ensure_malloc(size_t bytes, cudaStream_t stream, std::string* err) {
  void* nb = nullptr;
  cudaError_t e = cudaMalloc(&nb, bytes);
  if (e != cudaSuccess) {
    *err = std::string("cudaMalloc: ") + cudaGetErrorString(e);
    cudaGetLastError();
    return false;
  }
  if (base_) {
    cudaMemcpyAsync(nb, base_, mapped_, cudaMemcpyDeviceToDevice, stream);
    ++copies_;
  }
  cudaMemsetAsync((char*)nb + mapped_, 0, bytes - mapped_, stream);
  if (base_) {
    cudaStreamSynchronize(stream);
    cudaFree(base_);
  }
  base_ = nb;
  mapped_ = bytes;
  return true;
}

}  // namespace cslam
