// TEST INFRASTRUCTURE ONLY (oracle/): pass-through shl_array<T> so the restated generated
// procedures (pagerank.cc / hop_dist.cc in this directory) keep the reference's call shape
// (src/backend_cpp/gm_cpp_gen.cc:687-775 alloc/copy_from, :1468-1503 copy_back/print_crc).
// Plain host arrays; no placement policy; CRC printed as "n.a." (the polynomial lives in the
// absent libshoal contrib/pycrc, reference Makefile:98-100).
#ifndef ORACLE_REF_GLUE_SHL_ARRAYS_HPP
#define ORACLE_REF_GLUE_SHL_ARRAYS_HPP
#include <stdio.h>
#include <string.h>
#include <stddef.h>

template <class T>
class shl_array {
  public:
    shl_array(size_t n, const char* nm) : size(n), name(nm), data(NULL) {}
    virtual ~shl_array() { delete[] data; }
    virtual void alloc() { if (!data) data = new T[size ? size : 1]; }
    virtual T* get_array() { return data; }
    virtual void copy_from(T* src) { if (src) memcpy(data, src, size * sizeof(T)); }
    virtual void copy_back(T* dst) { if (dst) memcpy(dst, data, size * sizeof(T)); }
    virtual void print_crc() { printf("CRC %s n.a.\n", name); }
    void set_dynamic(bool) {}
    void set_used(bool) {}
    size_t get_size() { return size; }
  protected:
    size_t size;
    const char* name;
    T* data;
};

template <class T>
struct arr_thread_ptr { T* ptr1; T* ptr2; T* rep_ptr; };

template <class T>
shl_array<T>* shl__malloc_array(size_t n, const char* name, bool /*is_ro*/, bool /*is_dynamic*/,
                                bool /*is_used*/, bool /*is_graph*/, bool /*is_indexed*/,
                                bool /*do_init*/) {
    return new shl_array<T>(n, name);
}

template <class T>
void shl__wr_rep_ptr_thread_init(shl_array<T>* a, arr_thread_ptr<T>* p) {
    p->ptr1 = p->ptr2 = p->rep_ptr = a->get_array();
}
#endif
