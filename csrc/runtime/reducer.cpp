// b200ml — gradient bucket manager (host side of the data-parallel engine).
//
// Role of torch's c10d Reducer in the reference (`DistributedDataParallel(model)`, distributed_cnn.py:156,
// distributed_lstm.py:160, distributed_multilayer_perceptron.py:106): group parameters into buckets and fire one
// collective per bucket as soon as the bucket's last gradient exists.  Here the parameters already live in ONE flat
// buffer (parallel/flat.py), so a bucket is a contiguous range of it; the gradient kernels write straight into their
// slots and report "parameter i is ready"; the first report that completes a bucket returns its index and the engine
// launches the fused all-reduce + optimizer kernel on that range.  Buckets are cut at `cap` elements; parameters of
// cap/4 or more start a bucket of their own (see below).  Parameters are partitioned into groups (e.g.
// "weights whose fp32 master may stay stale on non-owner ranks" vs the rest) and buckets never straddle a group.
// Reports arrive from the autograd worker thread and the main thread, hence the mutex.
#include <mutex>
#include <vector>

#include "../../include/b200ml.h"
#include "../common/host.h"

struct b200ml_reducer {
    struct Bucket {
        long long start, numel;
        int first_param, n_params, group;
    };
    std::vector<Bucket> buckets;
    std::vector<int> bucket_of;   // parameter -> bucket
    std::vector<int> pending;     // per bucket: parameters that have not reported in this step
    std::vector<char> reported;   // per parameter
    std::mutex mu;
};

extern "C" {

b200ml_reducer* b200ml_reducer_create(int n_params, const long long* offsets, int n_groups, const int* group_ends,
                                      long long cap_elems) {
    using namespace b200;
    if (n_params <= 0 || !offsets || n_groups <= 0 || !group_ends || cap_elems <= 0) {
        set_error("reducer_create: bad arguments");
        return nullptr;
    }
    auto* r = new b200ml_reducer();
    r->bucket_of.assign(n_params, -1);
    int lo = 0;
    for (int g = 0; g < n_groups; ++g) {
        const int hi = group_ends[g];
        if (hi < lo || hi > n_params) {
            set_error("reducer_create: group boundaries must be non-decreasing and <= n_params");
            delete r;
            return nullptr;
        }
        // Bucket boundaries.  Ordinary parameters are packed until the bucket reaches `cap`.  A parameter of a quarter
        // of the cap or more ("big": an embedding table, the output projection) never joins the parameters registered
        // BEFORE it - their gradients appear at another time of backward, and a shared bucket would hold the big one's
        // all-reduce / optimizer launch back - but it takes the small parameters registered right AFTER it along, up
        // to cap/2 of them: for an embedding table those are the weights of the layer it feeds, whose gradients become
        // ready just before the table's, at the end of backward, so one launch (one pair of cross-GPU barriers)
        // serves both instead of two launches queued behind each other.
        long long start = lo < hi ? offsets[lo] : 0;
        int first = lo;
        long long big_end = -1;                       // end offset of the big parameter of the open bucket, if any
        auto is_big = [&](int i) { return (offsets[i + 1] - offsets[i]) * 4 >= cap_elems; };
        for (int i = lo; i < hi; ++i) {
            if (is_big(i)) big_end = offsets[i + 1];
            const long long end = offsets[i + 1];
            bool close = i + 1 == hi;
            if (!close) {
                if (is_big(i + 1)) close = true;
                else if (big_end >= 0) close = offsets[i + 2] - big_end > cap_elems / 2;
                else close = end - start >= cap_elems;
            }
            if (close) {
                r->buckets.push_back({start, end - start, first, i + 1 - first, g});
                for (int p = first; p <= i; ++p) r->bucket_of[p] = (int)r->buckets.size() - 1;
                start = end;
                first = i + 1;
                big_end = -1;
            }
        }
        lo = hi;
    }
    r->pending.assign(r->buckets.size(), 0);
    r->reported.assign(n_params, 0);
    return r;
}

void b200ml_reducer_destroy(b200ml_reducer* r) { delete r; }

int b200ml_reducer_num_buckets(const b200ml_reducer* r) { return r ? (int)r->buckets.size() : 0; }

int b200ml_reducer_bucket(const b200ml_reducer* r, int b, long long* start, long long* numel, int* first_param,
                          int* n_params, int* group) {
    using namespace b200;
    B200_REQUIRE(r && b >= 0 && b < (int)r->buckets.size(), "reducer_bucket: index %d out of range", b);
    const auto& k = r->buckets[b];
    if (start) *start = k.start;
    if (numel) *numel = k.numel;
    if (first_param) *first_param = k.first_param;
    if (n_params) *n_params = k.n_params;
    if (group) *group = k.group;
    return 0;
}

int b200ml_reducer_bucket_of(const b200ml_reducer* r, int param) {
    return (r && param >= 0 && param < (int)r->bucket_of.size()) ? r->bucket_of[param] : -1;
}

void b200ml_reducer_begin_step(b200ml_reducer* r) {
    if (!r) return;
    std::lock_guard<std::mutex> lock(r->mu);
    for (size_t b = 0; b < r->buckets.size(); ++b) r->pending[b] = r->buckets[b].n_params;
    std::fill(r->reported.begin(), r->reported.end(), 0);
}

// Returns the bucket that this report completed, -1 if its bucket still waits for other parameters (or the parameter
// reported twice in this step: gradient accumulation over a re-used weight), -2 on a bad index.
int b200ml_reducer_mark_ready(b200ml_reducer* r, int param) {
    if (!r || param < 0 || param >= (int)r->bucket_of.size()) return -2;
    std::lock_guard<std::mutex> lock(r->mu);
    if (r->reported[param]) return -1;
    r->reported[param] = 1;
    const int b = r->bucket_of[param];
    return --r->pending[b] == 0 ? b : -1;
}

// Buckets that still wait at the end of backward (parameters unused in this forward): marks them done and lists them.
int b200ml_reducer_flush(b200ml_reducer* r, int* out, int cap) {
    if (!r) return 0;
    std::lock_guard<std::mutex> lock(r->mu);
    int n = 0;
    for (size_t b = 0; b < r->buckets.size(); ++b)
        if (r->pending[b] > 0) {
            r->pending[b] = 0;
            if (out && n < cap) out[n] = (int)b;
            ++n;
        }
    return n;
}

}  // extern "C"
