// hemi::Array<T> over the C ABI of libsxmc_b200 -- the buffer handle type of the whole sxmc
// class API (reference: contrib/hemi/hemi/array.h:21-267; method names and validity-flag
// semantics follow it so src/*.cpp callers compile unchanged).  The host half is pinned
// memory (sxmc_host_alloc), the device half HBM (sxmc_buf_alloc); a copy happens only when
// the side asked for is stale, exactly as the reference's lazy host<->device mirroring.
#ifndef SXMC_B200_HEMI_ARRAY_H
#define SXMC_B200_HEMI_ARRAY_H

#include <cstddef>
#include <cstring>
#include <stdexcept>
#include <string>

#include "sxmc_b200.h"

namespace hemi {

enum Location { host = 0, device = 1 };

inline void abi_check(int rc, const char* what) {
  if (rc != SXMC_OK) throw std::runtime_error(std::string(what) + ": " + sxmc_last_error());
}

template <typename T>
class Array {
 public:
  // array.h:25-37 (usePinned is honoured as "always": pinned memory is what the async copies want)
  Array(size_t n, bool /*usePinned*/ = true)
      : n_(n), h_(nullptr), d_(nullptr), foreign_(false), h_valid_(false), d_valid_(false) {}
  // array.h:40-52: wrap caller-owned host memory
  Array(T* host_mem, size_t n)
      : n_(n), h_(host_mem), d_(nullptr), foreign_(true), h_valid_(true), d_valid_(false) {}
  ~Array() {
    if (h_ && !foreign_) sxmc_host_free(h_);
    if (d_) sxmc_buf_free(d_);
  }
  Array(const Array&) = delete;
  Array& operator=(const Array&) = delete;

  size_t size() const { return n_; }

  void copyFromHost(const T* other, size_t n) {       // array.h:64-72
    need_host();
    std::memcpy(h_, other, sizeof(T) * clamp(n));
    h_valid_ = true;
    d_valid_ = false;
  }
  void copyToHost(T* other, size_t n) const {         // array.h:75-80
    std::memcpy(other, readOnlyHostPtr(), sizeof(T) * clamp(n));
  }
  void copyFromDevice(const T* other_dev, size_t n) { // array.h:83-92 (device-to-device through the host mirror)
    need_host();
    abi_check(sxmc_buf_download(h_, other_dev, sizeof(T) * clamp(n)), "hemi::Array::copyFromDevice");
    h_valid_ = true;
    d_valid_ = false;
  }
  void copyToDevice(T* other_dev, size_t n) {         // array.h:95-103
    abi_check(sxmc_buf_upload(other_dev, readOnlyHostPtr(), sizeof(T) * clamp(n)), "hemi::Array::copyToDevice");
  }

  T* ptr(Location loc = host) { return loc == host ? hostPtr() : devicePtr(); }
  const T* readOnlyPtr(Location loc = host) const { return loc == host ? readOnlyHostPtr() : readOnlyDevicePtr(); }
  T* writeOnlyPtr(Location loc = host) { return loc == host ? writeOnlyHostPtr() : writeOnlyDevicePtr(); }

  // read-write access invalidates the other side (array.h:112-128)
  T* hostPtr() {
    sync_host();
    d_valid_ = false;
    return h_;
  }
  T* devicePtr() {
    sync_device();
    h_valid_ = false;
    return d_;
  }
  // read-only access leaves both sides valid (array.h:137-150)
  const T* readOnlyHostPtr() const {
    const_cast<Array*>(this)->sync_host();
    return h_;
  }
  const T* readOnlyDevicePtr() const {
    const_cast<Array*>(this)->sync_device();
    return d_;
  }
  // write-only access skips the copy (array.h:159-174)
  T* writeOnlyHostPtr() {
    need_host();
    h_valid_ = true;
    d_valid_ = false;
    return h_;
  }
  T* writeOnlyDevicePtr() {
    need_device();
    d_valid_ = true;
    h_valid_ = false;
    return d_;
  }

 private:
  size_t clamp(size_t n) const { return n < n_ ? n : n_; }
  void need_host() {
    if (!h_) {
      void* p = nullptr;
      abi_check(sxmc_host_alloc(sizeof(T) * (n_ ? n_ : 1), &p), "hemi::Array host allocation");
      h_ = static_cast<T*>(p);
      std::memset(h_, 0, sizeof(T) * n_);
    }
  }
  void need_device() {
    if (!d_) {
      void* p = nullptr;
      abi_check(sxmc_buf_alloc(sizeof(T) * (n_ ? n_ : 1), &p), "hemi::Array device allocation");
      d_ = static_cast<T*>(p);
    }
  }
  void sync_host() {
    need_host();
    if (!h_valid_ && d_valid_)
      abi_check(sxmc_buf_download(h_, d_, sizeof(T) * n_), "hemi::Array device->host");
    h_valid_ = true;
  }
  void sync_device() {
    need_device();
    if (!d_valid_) {
      if (h_valid_) abi_check(sxmc_buf_upload(d_, h_, sizeof(T) * n_), "hemi::Array host->device");
      else abi_check(sxmc_buf_zero(d_, sizeof(T) * n_), "hemi::Array zero");
    }
    d_valid_ = true;
  }

  size_t n_;
  T* h_;
  T* d_;
  bool foreign_;
  bool h_valid_, d_valid_;
};

}  // namespace hemi

#endif
