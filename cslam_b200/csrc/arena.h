// Growable HBM arena for the descriptor database.
//
// The reference grows its DB with torch.concat (cslam/nns_matching.py:38): allocate 2x, copy the
// whole DB device-to-device, free — 1 GB read + 2 GB written at the 1M -> 2M step of a 256-d DB.
// Here the DB lives in a reserved virtual address range and physical chunks are mapped behind it
// as capacity doubles (CUDA VMM: cuMemAddressReserve / cuMemCreate / cuMemMap), so growth never
// copies a row.  If the reserved range itself runs out, the SAME physical chunks are re-mapped into
// a larger range (still no copy).  The driver API is reached through cudaGetDriverEntryPoint, so the
// library has no link-time dependency on libcuda and loads on a machine without a driver.
// CSLAM_NNS_NO_VMM=1 (or a driver without VMM support) selects a cudaMalloc + copy fallback.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

namespace cslam {

struct DriverApi {
  bool ok = false;
  CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*memAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                        unsigned long long) = nullptr;
  CUresult (*memRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle,
                     unsigned long long) = nullptr;
  CUresult (*memUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*memGetAllocationGranularity)(size_t*, const CUmemAllocationProp*,
                                          CUmemAllocationGranularity_flags) = nullptr;
  void* tensorMapEncodeTiled = nullptr;
};
const DriverApi& driver_api();

class Arena {
 public:
  Arena() = default;
  ~Arena();
  Arena(const Arena&) = delete;
  Arena& operator=(const Arena&) = delete;

  // Make at least `bytes` of zero-initialised storage addressable at base().  Existing contents are
  // preserved.  Returns false and sets err on failure.  May change base() (never copies when VMM is
  // in use); the caller must have synchronised outstanding work if it does (it does so itself).
  bool ensure(int device, size_t bytes, cudaStream_t stream, std::string* err);
  void* base() const { return base_; }
  size_t mapped() const { return mapped_; }
  bool uses_vmm() const { return vmm_; }
  int chunks() const { return (int)chunks_.size(); }
  long long full_copies() const { return copies_; }
  void release();

 private:
  struct Chunk {
    CUmemGenericAllocationHandle handle;
    size_t size;
  };
  bool ensure_vmm(int device, size_t bytes, cudaStream_t stream, std::string* err);
  bool ensure_malloc(size_t bytes, cudaStream_t stream, std::string* err);

  void* base_ = nullptr;
  size_t mapped_ = 0;    // bytes usable at base_
  size_t reserved_ = 0;  // VA bytes reserved at base_ (VMM only)
  size_t gran_ = 0;
  bool vmm_ = false;
  bool decided_ = false;
  long long copies_ = 0;
  std::vector<Chunk> chunks_;
};

}  // namespace cslam
