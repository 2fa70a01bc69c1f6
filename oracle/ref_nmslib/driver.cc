// driver.cc -- thin C-ABI driver over the REFERENCE's own vendored NMSLIB sources.
//
// TEST INFRASTRUCTURE ONLY (see oracle/knn_oracle.c header).  Built by oracle/Makefile into
// oracle/_ref/libnmslib_ref.so from the sources where they lie under
// /root/reference/matchinglib_poselib/source/matchinglib/similarity_search (include path
// only; nothing is copied).  The only part restated here is the 260-line OpenCV-typed
// wrapper nmslibMatching<dist_t> (MP/matchinglib/include/nmslib/nmslib_matchers.h:159-424),
// which needs OpenCV C++ headers that this image does not have: packing (:212-231,277-296),
// K=2 (:180), q % ThreadQty striping (:104-130), pop + sort-by-distance (:352-377).
// Everything below Space/Index/KNNQuery is the reference's compiled code.
#include <cstdint>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>
#include <algorithm>

#include "space/space_bit_hamming.h"
#include "space/space_lp.h"
#include "method/seqsearch.h"
#include "knnquery.h"
#include "knnqueue.h"
#include "object.h"
#include "logging.h"
#include "params.h"

using namespace similarity;

template <typename dist_t>
static void search_all(Space<dist_t>& space, const ObjectVector& data, const ObjectVector& queries,
                       int threads, int32_t* idx, dist_t* dist)
{
    SeqSearch<dist_t> index(space, data);
    AnyParams empty;
    index.CreateIndex(empty);
    index.SetQueryTimeParams(empty);
    space.SetQueryPhase();
    const unsigned K = 2;                       // nmslib_matchers.h:180
    auto worker = [&](unsigned part) {
        for (size_t q = 0; q < queries.size(); ++q) {
            if ((q % (size_t)threads) != part) continue;      // :112-113
            KNNQuery<dist_t> query(space, queries[q], K, 0.0);
            index.Search(&query, -1);
            std::unique_ptr<KNNQueue<dist_t>> res(query.Result()->Clone());
            // :352-377  pop the max-heap, then sort ascending by distance
            std::vector<std::pair<dist_t, int32_t>> v;
            while (!res->Empty()) { v.emplace_back(res->TopDistance(), (int32_t)res->TopObject()->id()); res->Pop(); }
            std::stable_sort(v.begin(), v.end(), [](const std::pair<dist_t,int32_t>& a, const std::pair<dist_t,int32_t>& b) { return a.first < b.first; });
            for (size_t j = 0; j < 2; ++j) {
                idx[q * 2 + j]  = j < v.size() ? v[j].second : -1;
                dist[q * 2 + j] = j < v.size() ? v[j].first : (dist_t)0;
            }
        }
    };
    if (threads > 1) {
        std::vector<std::thread> th;
        for (int p = 0; p < threads; ++p) th.emplace_back(worker, (unsigned)p);
        for (auto& t : th) t.join();
    } else worker(0);
}

static void pack_rows(const uint8_t* d, int n, size_t step, int nbytes, ObjectVector& out)
{
    std::vector<int> vec;
    for (int i = 0; i < n; ++i) {
        const uint8_t* r = d + (size_t)i * step;
        vec.clear();
        for (int j = 0; j < nbytes - 1; j += 2) vec.push_back(((int)r[j] << 8) | (int)r[j + 1]);   // :212-225
        if (nbytes & 1) vec.push_back((int)r[nbytes - 1]);
        out.push_back(new Object(i, -1, vec.size() * sizeof(int), vec.data()));                    // :138-140
    }
}

extern "C" int ref_nmslib_knn2_hamming(const uint8_t* q, int nq, size_t q_step, const uint8_t* t, int nt, size_t t_step,
                                       int nbytes, int threads, int32_t* idx, int32_t* dist)
{
    InitializeLogger(LIB_LOGNONE, NULL);   // nmslib_matchers.h:191 initLibrary(LIB_LOGNONE) minus the space/method factories (objects are built directly)
    ObjectVector data, queries;
    pack_rows(t, nt, t_step, nbytes, data);
    pack_rows(q, nq, q_step, nbytes, queries);
    SpaceBitHamming space;
    search_all<int>(space, data, queries, threads, idx, dist);
    for (auto o : data) delete o;
    for (auto o : queries) delete o;
    return 0;
}

extern "C" int ref_nmslib_knn2_l2(const float* q, int nq, size_t q_step, const float* t, int nt, size_t t_step,
                                  int dim, int threads, int32_t* idx, float* dist)
{
    InitializeLogger(LIB_LOGNONE, NULL);   // nmslib_matchers.h:191 initLibrary(LIB_LOGNONE) minus the space/method factories (objects are built directly)
    ObjectVector data, queries;
    for (int i = 0; i < nt; ++i) data.push_back(new Object(i, -1, dim * sizeof(float), (const char*)t + (size_t)i * t_step));
    for (int i = 0; i < nq; ++i) queries.push_back(new Object(i, -1, dim * sizeof(float), (const char*)q + (size_t)i * q_step));
    SpaceLp<float> space(2);
    search_all<float>(space, data, queries, threads, idx, dist);
    for (auto o : data) delete o;
    for (auto o : queries) delete o;
    return 0;
}
