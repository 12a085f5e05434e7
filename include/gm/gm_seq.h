// gm_seq<T>: the ordered collection type behind Green-Marl's Node_Sequence / Edge_Sequence (reference
// apps/output_cpp/gm_graph/inc/gm_seq.h:31-147: gm_seq, Seq_Iterator, typedefs gm_node_seq / gm_edge_seq).
// Same public API, so drivers (bc_main.cc:21-38: `new gm_node_seq()`, push_back) and generated code
// (`gm_node_seq::seq_iter I = Seeds.prepare_seq_iteration(); while (I.has_next()) { node_t s = I.get_next(); ...`)
// compile unchanged.  Host-only container: a deque plus per-thread staging buffers for the *_par calls.
#ifndef GM_SEQ_H
#define GM_SEQ_H
#include <assert.h>
#include <deque>
#include <vector>
#include "gm_graph_typedef.h"
#include "gm_runtime.h"

template <typename IterType, typename T>
class Seq_Iterator {
  public:
    Seq_Iterator(IterType first, IterType last) : cur(first), stop(last) {}
    inline bool has_next() { return cur != stop; }
    inline T get_next() { T v = *cur; ++cur; return v; }

  private:
    IterType cur, stop;
};

template <typename T>
class gm_seq {
  public:
    gm_seq() : front_stage(stage_count(gm_rt_get_num_threads())), back_stage(front_stage.size()) {}
    explicit gm_seq(int max_threads) : front_stage(stage_count(max_threads)), back_stage(front_stage.size()) {}
    virtual ~gm_seq() {}

    void push_back(T e) { items.push_back(e); }
    void push_front(T e) { items.push_front(e); }
    T pop_back() { T e = items.back(); items.pop_back(); return e; }
    T pop_front() { T e = items.front(); items.pop_front(); return e; }
    void clear() { items.clear(); }
    int get_size() { return (int)items.size(); }

    // parallel insertion: staged per thread, made visible by merge() (front stages before the body, back stages after)
    void push_back_par(T e, int tid) { back_stage[tid].push_back(e); }
    void push_front_par(T e, int tid) { front_stage[tid].push_front(e); }
    void merge() {
        for (size_t t = 0; t < front_stage.size(); t++) {
            items.insert(items.begin(), front_stage[t].begin(), front_stage[t].end());
            items.insert(items.end(), back_stage[t].begin(), back_stage[t].end());
            front_stage[t].clear(); back_stage[t].clear();
        }
    }

    typedef Seq_Iterator<typename std::deque<T>::iterator, T> seq_iter;
    typedef Seq_Iterator<typename std::deque<T>::reverse_iterator, T> rev_iter;
    typedef seq_iter par_iter;
    seq_iter prepare_seq_iteration() { return seq_iter(items.begin(), items.end()); }
    rev_iter prepare_rev_iteration() { return rev_iter(items.rbegin(), items.rend()); }

  private:
    static size_t stage_count(int threads) { return (size_t)(threads > 0 ? threads : 1); }
    std::deque<T> items;
    std::vector<std::deque<T> > front_stage, back_stage;
};

typedef gm_seq<node_t> gm_node_seq;
typedef gm_seq<edge_t> gm_edge_seq;
#endif
