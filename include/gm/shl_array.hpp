// shl_array<T>: Shoal's array abstraction, re-targeted to HBM.  Surface inferred from the reference's
// call sites (libshoal itself is not in the tree): src/backend_cpp/gm_cpp_gen.cc:687-775 (malloc_array,
// alloc, copy_from), :1468-1503 (copy_back, print_crc), src/inc/gm_backend_cpp.h:671-704 (get_array).
//
// Placement policy: where Shoal chooses between replicated / partitioned / distributed NUMA pages, the
// array here lives in the HBM of the current device; get_array() returns the DEVICE pointer that kernels
// index (host code must not dereference it).  copy_from / copy_back are the H2D / D2H copies.
#ifndef SHL_ARRAY_HPP_
#define SHL_ARRAY_HPP_
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "gm_cuda.h"

extern "C" uint16_t shl__crc16(const void* data, size_t bytes);

template <class T>
class shl_array {
  public:
    shl_array(size_t n, const char* nm, bool ro = false, bool dyn = false, bool used = true, bool graph = false, bool indexed = false)
        : size(n), name(nm), dev(NULL), is_ro(ro), is_dynamic(dyn), is_used(used), is_graph(graph), is_indexed(indexed) {}
    virtual ~shl_array() { if (dev) gmc_free(dev); }

    virtual void alloc() {
        if (dev || !is_used) return;
        if (gmc_malloc((void**)&dev, sizeof(T) * (size ? size : 1)) != GMC_OK) die("alloc");
        print_options();
    }
    virtual T* get_array() { return dev; }                       // device pointer
    virtual void copy_from(T* src) {
        if (!dev || !src) return;
        if (gmc_h2d(dev, src, sizeof(T) * size, NULL) != GMC_OK || gmc_sync(NULL) != GMC_OK) die("copy_from");
    }
    virtual void copy_back(T* dst) {
        if (!dev || !dst || is_ro) return;
        if (gmc_d2h(dst, dev, sizeof(T) * size, NULL) != GMC_OK) die("copy_back");
    }
    // "CRC <name> 0x%04x" over the array contents (stdout contract of scripts/extract_result.py:162-167).
    // CRC-16/CCITT-FALSE; the polynomial libshoal uses is not recoverable from the reference tree.
    virtual void print_crc() {
        if (!dev || size == 0) { printf("CRC %s n.a.\n", name); return; }
        T* tmp = (T*)malloc(sizeof(T) * size);
        if (!tmp || gmc_d2h(tmp, dev, sizeof(T) * size, NULL) != GMC_OK) die("print_crc");
        printf("CRC %s 0x%04x\n", name, (unsigned)shl__crc16(tmp, sizeof(T) * size));
        free(tmp);
    }
    void set_dynamic(bool v) { is_dynamic = v; }
    void set_used(bool v) { is_used = v; }
    size_t get_size() { return size; }
    const char* get_name() { return name; }

  protected:
    void print_options() {
        shl__print_array_conf(name, size, sizeof(T), is_used ? 1 : 0, 0, 0);
    }
    void die(const char* what) {
        fprintf(stderr, "shl_array[%s]::%s failed: %s\n", name, what, gmc_last_error());
        exit(EXIT_FAILURE);
    }
    size_t size;
    const char* name;
    T* dev;
    bool is_ro, is_dynamic, is_used, is_graph, is_indexed;
};

// write-replicated arrays (reference gm_cpp_gen.cc:361-383,712-731): one HBM copy per device is the
// multi-GPU equivalent; on one device it degenerates to the plain array.
template <class T>
class shl_array_wr_rep : public shl_array<T> {
  public:
    shl_array_wr_rep(size_t n, const char* nm, int (*rep_id)(void)) : shl_array<T>(n, nm), get_rep_id(rep_id) {}
  private:
    int (*get_rep_id)(void);
};
#endif
