// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// The entry points of oracle/mih.cpp around THE REFERENCE'S OWN HASHING::MultiIndexHashing: src/Hashing.cc is compiled
// whole and unmodified next to this file (oracle/Makefile, target ref), against the MapPoint stand-in of
// oracle/ref_mih_stub.hpp.  Entries are MapPoint stand-ins owned by the handle, one per entry id (mnId); queries report
// the ids of the MapPoint* the reference returns.  A separate library (libgforb_ref_mih.so): the stand-in MapPoint of
// the matcher build (oracle/ref_stub.hpp) is a different class of the same name.
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "Hashing.h"                 // the reference's own header (-I/root/reference/include)

using ORB_SLAM2::MapPoint;

// The reference's constructor prints through std::cout (src/Hashing.cc:513-518).  This library is built by a newer g++
// than the libstdc++ a host process may already have loaded (numpy brings the system's): new <iostream> headers leave the
// stream set-up to the runtime, the older runtime leaves it to an ios_base::Init object in some translation unit.
static std::ios_base::Init gforb_ios_init;

struct orc_mih {
    HASHING::MultiIndexHashing* mih;
    int table_num;
    std::map<int, MapPoint*> points;
    MapPoint* point(int id) {
        auto it = points.find(id);
        if (it != points.end()) return it->second;
        MapPoint* p = new MapPoint;
        p->mnId = (long unsigned)id;
        p->desc_ = cv::Mat(1, 32, CV_8UC1);
        return points[id] = p;
    }
};

extern "C" {

orc_mih* orc_mih_create(int feature_length, int table_num) {
    if (feature_length != 256 || table_num < 1 || table_num > 32 || feature_length / table_num > 20) return nullptr;
    orc_mih* h = new orc_mih;
    h->table_num = table_num;
    h->mih = new HASHING::MultiIndexHashing((unsigned)feature_length, (unsigned)table_num);
    return h;
}
void orc_mih_destroy(orc_mih* h) {
    if (!h) return;
    for (auto& kv : h->points) delete kv.second;
    delete h->mih;
    delete h;
}
int orc_mih_substring_len(const orc_mih* h) { return (int)h->mih->getSubstringLen(); }

// hash() is `inline` in the class and defined in Hashing.cc, so it is not callable from here; the keys are observable
// through insert + query and are pinned that way (tests/test_oracle_mih.py)
void orc_mih_insert(orc_mih* h, const uint8_t* desc, const int32_t* entry, int n, uint32_t* hashed, uint8_t* success) {
    for (int i = 0; i < n; ++i) {
        MapPoint* p = h->point(entry[i]);
        std::memcpy(p->desc_.ptr(0), desc + 32 * (size_t)i, 32);
        if (hashed) for (int t = 0; t < 32; ++t) p->mvbHashed[t] = (hashed[i] >> t) & 1u;
        const bool ok = h->mih->insert(p);
        if (hashed) { uint32_t f = 0; for (int t = 0; t < 32; ++t) f |= (uint32_t)p->mvbHashed[t] << t; hashed[i] = f; }
        if (success) success[i] = ok;
    }
}

int orc_mih_query(orc_mih* h, const uint8_t* qdesc, int m, int nactive, int nretrieve, int32_t* cand, int stride, int32_t* count) {
    int worst = 0;
    std::vector<HASHING::Entry> c;
    for (int q = 0; q < m; ++q) {
        unsigned d[8];
        std::memcpy(d, qdesc + 32 * (size_t)q, 32);
        h->mih->query(d, c, (unsigned)nactive, (unsigned)nretrieve);
        const int k = (int)c.size();
        for (int i = 0; i < k && i < stride; ++i) cand[(size_t)q * stride + i] = (int32_t)c[i]->mnId;
        count[q] = k;
        worst = k > worst ? k : worst;
    }
    return worst;
}

int orc_mih_query2(orc_mih* h, const uint8_t* qdesc, int m, const int32_t* nfeatures, int32_t* cand, int32_t* count,
                   int32_t* addi, int32_t* addi_count, int stride) {
    int worst = 0;
    std::vector<HASHING::Entry> c, a;
    std::vector<int> nf(nfeatures, nfeatures + h->table_num);
    for (int q = 0; q < m; ++q) {
        unsigned d[8];
        std::memcpy(d, qdesc + 32 * (size_t)q, 32);
        h->mih->query(d, c, a, nf);
        const int k = (int)c.size(), ka = (int)a.size();
        for (int i = 0; i < k && i < stride; ++i) cand[(size_t)q * stride + i] = (int32_t)c[i]->mnId;
        for (int i = 0; i < ka && i < stride; ++i) addi[(size_t)q * stride + i] = (int32_t)a[i]->mnId;
        count[q] = k; addi_count[q] = ka;
        worst = std::max(worst, std::max(k, ka));
    }
    return worst;
}

int orc_mih_update_dynamics(orc_mih* h, double e) { return h->mih->updateDynamics(e); }

}  // extern "C"
