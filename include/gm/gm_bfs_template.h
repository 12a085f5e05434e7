// gm_bfs_template: level-synchronous BFS with user callbacks, API-compatible with the reference class
// template (apps/output_cpp/gm_graph/inc/gm_bfs_template.h:14-754) and re-targeted to the device.
//
// Reference behaviour being kept: prepare(root, nthreads); do_bfs_forward() visits every vertex reachable
// from root level by level, calling visit_fw(t) while get_curr_level() is t's level and
// do_end_of_level_fw() after each level; do_bfs_reverse() walks the levels back calling visit_rv(t) /
// do_end_of_level_rv(); is_down_edge(e) is true for an edge t->u with level(u) == level(t)+1 (:483-524,
// :581-622) unless t's level was expanded bottom-up, where the reference marks nothing (:215-259,:689-708; a save_child
// instance reaches those states only through RD -> RRD, :398-400); check_navigator(u, e) filters edges (:488-492);
// levels are level_t with -2 = unvisited (:725).
//
// Device formulation: the TRAVERSAL (the reference's iterate_neighbor_que/rd + check_parent_rrd state
// machine, :69-271,:352-421) runs on the GPU in one call (gmc_bfs_levels*, include/gm_cuda.h) over the
// graph's HBM mirror.  It returns visited_level[], the per-level counts and the level-ordered vertex
// list (the reference's global_vector / level_queue_begin, :423-447).  The user's virtuals are host C++,
// so they are then replayed on the host, level by level, over those queues with the same
// curr_level / get_level() view the reference offers.  The order of visits INSIDE a level is unspecified
// in the reference too (dynamic OpenMP schedule, :112,:139).  A navigator is evaluated once per edge into
// an edge mask before the traversal (the reference evaluates it lazily, on first sight of an unvisited
// neighbour; Green-Marl navigators are pure filter expressions, so both see the same value).
#ifndef GM_BFS_TEMPLATE_H
#define GM_BFS_TEMPLATE_H
#include <omp.h>
#include <string.h>
#include <map>
#include <set>
#include <vector>

#include "gm_graph.h"
#include "gm_atomic_wrapper.h"
#include "gm_bitmap.h"
#include "gm_cuda.h"

template <typename level_t, bool use_multithread, bool has_navigator, bool use_reverse_edge, bool save_child>
class gm_bfs_template {
  public:
    gm_bfs_template(gm_graph& _G) : G(_G) {
        visited_bitmap = NULL;
        visited_level = NULL;
        down_edge_array = NULL;
        is_finished = false;
        curr_level = 0;
        root = gm_graph::NIL_NODE;
        state = ST_SMALL;
        curr_count = next_count = 0;
        num_thread = 1;
    }

    virtual ~gm_bfs_template() {
        delete[] visited_bitmap;
        delete[] visited_level;
        delete[] down_edge_array;
    }

    void prepare(node_t root_node, int max_num_thread) {
        num_thread = use_multithread ? (max_num_thread > 0 ? max_num_thread : 1) : 1;
        is_finished = false;
        curr_level = 0;
        root = root_node;
        state = ST_SMALL;
        assert(root != gm_graph::NIL_NODE);
    }

    void do_bfs_forward() {
        const node_t n = G.num_nodes();
        const edge_t m = G.num_edges();
        if (use_reverse_edge && !G.has_reverse_edge()) G.make_reverse_edges();
        gmc_graph* dev = G.device_mirror();

        // optional edge filter: mask[e] = check_navigator(target(e), e) over the arrays being followed
        std::vector<unsigned char> mask;
        if (has_navigator) {
            mask.resize(m > 0 ? m : 1);
#pragma omp parallel for schedule(dynamic, 128) num_threads(num_thread)
            for (node_t t = 0; t < n; t++) {
                edge_t b, e;
                get_range(b, e, t);
                for (edge_t nx = b; nx < e; nx++) mask[nx] = check_navigator(get_node(nx), nx) ? 1 : 0;
            }
        }

        // traversal on the device.  The device entry point returns 16-bit levels whatever level_t is: the generator only
        // instantiates level_t = short (src/backend_cpp/gm_cpp_gen_bfs.cc:90); a wider level_t gets the same values widened, and
        // a traversal deeper than 32767 levels fails with an error (gmc_bfs_levels) instead of wrapping as the reference's short would.
        static_assert(sizeof(level_t) >= sizeof(int16_t), "gm_bfs_template: level_t narrower than the device's 16-bit levels");
        std::vector<int16_t> lev16(n > 0 ? n : 1);
        std::vector<int32_t> trace(LEVEL_CAP), counts(LEVEL_CAP);
        gmc_bfs_info info;
        const int flags = (use_reverse_edge ? GMC_BFS_REVERSE_EDGE : 0) | (save_child ? GMC_BFS_SAVE_CHILD : 0);
        int rc = gmc_bfs_levels_masked(dev, root, flags, has_navigator ? &mask[0] : NULL, &lev16[0],
                                       &trace[0], &counts[0], LEVEL_CAP, &info);
        if (rc != GMC_OK) die("gmc_bfs_levels");
        const int nlev = info.levels;
        if (nlev > LEVEL_CAP) die("more BFS levels than supported");
        global_vector.resize(n > 0 ? n : 1);
        std::vector<int32_t> lbegin(nlev + 1);
        rc = gmc_bfs_level_order(dev, &global_vector[0], &lbegin[0], nlev + 1);
        if (rc != GMC_OK) die("gmc_bfs_level_order");
        global_vector.resize(lbegin[nlev]);

        // host-visible state of the reference class
        delete[] visited_level;
        visited_level = new level_t[n > 0 ? n : 1];
        delete[] visited_bitmap;
        visited_bitmap = new unsigned char[(n + 7) / 8 + 1];
        memset(visited_bitmap, 0, (n + 7) / 8 + 1);
#pragma omp parallel for num_threads(num_thread)
        for (node_t i = 0; i < n; i++) visited_level[i] = (level_t)lev16[i];
        for (size_t i = 0; i < global_vector.size(); i++) _gm_set_bit(visited_bitmap, global_vector[i]);
        level_count.assign(counts.begin(), counts.begin() + nlev);
        level_queue_begin.assign(lbegin.begin(), lbegin.begin() + nlev);
        state_trace.assign(trace.begin(), trace.begin() + nlev);

        if (save_child) {   // edges t->u with level(u) == level(t)+1 that pass the navigator, of the levels expanded top-down:
                            // the bottom-up states mark none (check_parent_rrd, reference :689-708)
            delete[] down_edge_array;
            down_edge_array = new unsigned char[m > 0 ? m : 1];
#pragma omp parallel for schedule(dynamic, 128) num_threads(num_thread)
            for (node_t t = 0; t < n; t++) {
                edge_t b, e;
                get_range(b, e, t);
                const level_t lt = visited_level[t];
                const bool marking = lt != __INVALID_LEVEL && state_trace[lt] != ST_RRD && state_trace[lt] != ST_RR2Q;
                for (edge_t nx = b; nx < e; nx++) {
                    bool down = marking && visited_level[get_node(nx)] == (level_t)(lt + 1);
                    if (down && has_navigator) down = mask[nx] != 0;
                    down_edge_array[nx] = down ? 1 : 0;
                }
            }
        }

        // replay of the user's forward callbacks, level by level
        for (curr_level = 0;; curr_level++) {
            state = state_trace[curr_level];
            curr_count = level_count[curr_level];
            next_count = (curr_level + 1 < nlev) ? level_count[curr_level + 1] : 0;
            const node_t* q = &global_vector[level_queue_begin[curr_level]];
            if (use_multithread && curr_count >= THRESHOLD1) {
#pragma omp parallel for schedule(dynamic, 128) num_threads(num_thread)
                for (node_t i = 0; i < curr_count; i++) visit_fw(q[i]);
            } else {
                for (node_t i = 0; i < curr_count; i++) visit_fw(q[i]);
            }
            do_end_of_level_fw();
            if (curr_level + 1 >= nlev) break;
        }
        is_finished = true;   // curr_level stays at the last level, as in the reference (:262-270)
        state = ST_QUE;       // is_down_edge() reads the array form from here on
    }

    void do_bfs_reverse() {
        // only after do_bfs_forward (reference :273-312)
        level_t& level = curr_level;
        while (true) {
            const node_t count = level_count[level];
            const node_t* q = &global_vector[level_queue_begin[level]];
            if (use_multithread && count >= THRESHOLD1) {
#pragma omp parallel for schedule(dynamic, 128) num_threads(num_thread)
                for (node_t i = 0; i < count; i++) visit_rv(q[i]);
            } else {
                for (node_t i = 0; i < count; i++) visit_rv(q[i]);
            }
            do_end_of_level_rv();
            if (level == 0) break;
            level--;
        }
    }

    bool is_down_edge(edge_t idx) { return down_edge_array != NULL && down_edge_array[idx] != 0; }

    // ---- additions (not in the reference): what the device traversal did ----
    int num_levels() const { return (int)level_count.size(); }
    const std::vector<node_t>& get_level_count() const { return level_count; }
    const std::vector<int>& get_state_trace() const { return state_trace; }
    const level_t* get_levels() const { return visited_level; }

  protected:
    virtual void visit_fw(node_t t) = 0;
    virtual void visit_rv(node_t t) = 0;
    virtual bool check_navigator(node_t t, edge_t nx) = 0;
    virtual void do_end_of_level_fw() {}
    virtual void do_end_of_level_rv() {}

    node_t get_root() { return root; }
    level_t get_level(node_t t) { return visited_level[t]; }
    level_t get_curr_level() { return curr_level; }

    void get_range(edge_t& b, edge_t& e, node_t t) {
        if (use_reverse_edge) { b = G.r_begin[t]; e = G.r_begin[t + 1]; }
        else { b = G.begin[t]; e = G.begin[t + 1]; }
    }
    void get_range_rev(edge_t& b, edge_t& e, node_t t) {
        if (use_reverse_edge) { b = G.begin[t]; e = G.begin[t + 1]; }
        else { b = G.r_begin[t]; e = G.r_begin[t + 1]; }
    }
    node_t get_node(edge_t& nx) { return use_reverse_edge ? G.r_node_idx[nx] : G.node_idx[nx]; }
    node_t get_node_rev(edge_t& nx) { return use_reverse_edge ? G.node_idx[nx] : G.r_node_idx[nx]; }

    static const int ST_SMALL = 0;
    static const int ST_QUE = 1;
    static const int ST_Q2R = 2;
    static const int ST_RD = 3;
    static const int ST_R2Q = 4;
    static const int ST_RRD = 5;
    static const int ST_RR2Q = 6;
    static const int THRESHOLD1 = 128;
    static const int THRESHOLD2 = 1024;
    static const level_t __INVALID_LEVEL = -2;
    static const int LEVEL_CAP = 32768;

    int state;
    unsigned char* visited_bitmap;
    level_t* visited_level;
    bool is_finished;
    level_t curr_level;
    node_t root;
    gm_graph& G;
    node_t curr_count;
    node_t next_count;
    unsigned char* down_edge_array;
    std::vector<node_t> global_vector;
    std::vector<node_t> level_queue_begin;
    std::vector<node_t> level_count;
    std::vector<int> state_trace;
    int num_thread;

  private:
    void die(const char* what) {
        fprintf(stderr, "gm_bfs_template: %s: %s\n", what, gmc_last_error());
        exit(EXIT_FAILURE);   // no CPU fallback by design
    }
};
#endif
