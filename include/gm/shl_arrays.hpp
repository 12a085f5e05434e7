// Factory + helpers of the Shoal array API as the generated code calls them
// (src/backend_cpp/gm_cpp_gen.cc:181-227 COST macro, :687-775 malloc_array, src/inc/gm_backend_cpp.h:688-703).
#ifndef SHL_ARRAYS_HPP_
#define SHL_ARRAYS_HPP_
#include "shl_array.hpp"

template <class T>
struct arr_thread_ptr { T* ptr1; T* ptr2; T* rep_ptr; };

// Shoal's policy (partition if indexed, else replicate if read-only, else distribute) collapses to
// "resident in this device's HBM"; the flags are recorded for the Array[...] configuration line.
template <class T>
shl_array<T>* shl__malloc_array(size_t n, const char* name, bool is_ro, bool is_dynamic, bool is_used,
                                bool is_graph, bool is_indexed, bool /*do_init*/) {
    return new shl_array<T>(n, name, is_ro, is_dynamic, is_used, is_graph, is_indexed);
}

template <class T>
void shl__wr_rep_ptr_thread_init(shl_array<T>* a, arr_thread_ptr<T>* p) {
    p->ptr1 = p->ptr2 = p->rep_ptr = a->get_array();
}

template <class T>
size_t shl__estimate_size(size_t n, const char* /*name*/, bool /*ro*/, bool /*dyn*/, bool used, bool /*graph*/, bool /*indexed*/) {
    return used ? n * sizeof(T) : 0;
}
void shl__estimate_working_set_size(int n, ...);   // prints the HBM working set; sizes passed as size_t varargs
#endif
