// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// HASHING::MultiIndexHashing, the descriptor side (reference include/Hashing.h:171-216, src/Hashing.cc): substring keys
// (hash, :53-103), bucket insertion (insert :533-598 over Bucket::add :120-154 / front :319-484) and the two candidate
// queries (:699-846 and :849-950 over Bucket::back :250-279), restated over integer entry handles instead of MapPoint*:
// entry = MapPoint::mnId, the mvbHashed flags travel as a bit mask the caller keeps per entry.  The same entry points
// are exported by oracle/_ref/libgforb_ref_mih.so around the reference's own src/Hashing.cc compiled unmodified
// (oracle/ref_mih.cpp); tests/test_oracle_mih.py holds the two together.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <vector>

namespace {
constexpr int kMaxSupportedBitLength = 32;      // MAX_SUPPORTED_BIT_LENGTH, include/Hashing.h:56
constexpr int kNumCandidateFeatures = 20;       // NUM_CANDIDATE_FEATURES :60
constexpr int kMinBucketSize = 1, kMaxBucketSize = 20;   // :58-59
constexpr double kMinTimeDiff = 0.001;          // :83

struct Bucket {                                 // :107-150: newest entry first; front() = the entry added last
    std::deque<int> entries;
    int latest = -1;
    bool has_latest = false;
    void add(int e) { entries.push_front(e); latest = e; has_latest = true; }
    void back(std::vector<int>& dst, int n) const {          // :250-279
        dst.clear();
        if (n <= 0) return;                     // the reference returns on n == 0; a negative n is undefined there
        const int s = (int)entries.size();
        if (s == 0) return;
        if (n >= s) dst.insert(dst.end(), entries.begin(), entries.end());
        else dst.insert(dst.end(), entries.begin(), entries.begin() + n);
    }
};
}  // namespace

struct orc_mih {
    unsigned feature_length, hashtable_num, bits_per_substring;
    size_t bucket_num;
    int dynamics = 0;
    std::vector<std::vector<Bucket>> tables;

    void hash(const unsigned* pdesc, std::vector<size_t>& bucket_idx) const {      // :53-103
        if (bits_per_substring == 8) {
            const unsigned char* v = (const unsigned char*)pdesc;
            for (unsigned i = 0; i < hashtable_num; i++) bucket_idx[i] = v[i];
        } else if (bits_per_substring == 16) {
            const unsigned short* v = (const unsigned short*)pdesc;
            for (unsigned i = 0; i < hashtable_num; i++) bucket_idx[i] = v[i];
        } else {
            size_t mask[kMaxSupportedBitLength];
            for (int j = 0; j < kMaxSupportedBitLength; j++) {
                mask[j] = 0;
                for (int i = 0; i < j; i++) mask[j] += (size_t)1 << i;
            }
            for (unsigned i = 0; i < hashtable_num; i++) {
                size_t index = 0;
                const int si = (bits_per_substring * i) / kMaxSupportedBitLength;
                const int ei = (bits_per_substring * (i + 1)) / kMaxSupportedBitLength;
                const int sp = (bits_per_substring * i) % kMaxSupportedBitLength;
                const int ep = (bits_per_substring * (i + 1)) % kMaxSupportedBitLength;
                if (si == ei) index = (pdesc[si] >> sp) & mask[bits_per_substring];
                // si < ei with ep == 0 reads pdesc[ei] times mask[0] = 0: the word past the substring contributes nothing,
                // but the reference does index it; for the last table of a 256-bit descriptor that is pdesc[8] -- one word
                // past the row (its value is masked away)
                if (si < ei) index = ((pdesc[si] >> sp) & mask[kMaxSupportedBitLength - sp]) +
                                     (((ei < 8 ? pdesc[ei] : 0u) & mask[ep]) << (bits_per_substring - ep));
                bucket_idx[i] = index;
            }
        }
    }
};

extern "C" {

orc_mih* orc_mih_create(int feature_length, int table_num) {
    if (feature_length != 256 || table_num < 1 || table_num > 32) return nullptr;     // mvbHashed holds NUM_TOTAL_HASHTABLES = 32 flags
    orc_mih* h = new orc_mih;
    h->feature_length = feature_length; h->hashtable_num = table_num;
    h->bits_per_substring = feature_length / table_num;                                 // :492
    if (h->bits_per_substring > 20) { delete h; return nullptr; }     // 2^bits buckets per table are allocated up front (:504, :522-531)
    h->bucket_num = (size_t)std::pow(2, h->bits_per_substring);
    h->tables.assign(table_num, std::vector<Bucket>(h->bucket_num));
    return h;
}
void orc_mih_destroy(orc_mih* h) { delete h; }
int orc_mih_substring_len(const orc_mih* h) { return (int)h->bits_per_substring; }

void orc_mih_hash(const orc_mih* h, const uint8_t* desc, int n, uint32_t* keys) {
    std::vector<size_t> idx(h->hashtable_num);
    for (int i = 0; i < n; ++i) {
        unsigned d[8];
        std::memcpy(d, desc + 32 * (size_t)i, 32);
        h->hash(d, idx);
        for (unsigned t = 0; t < h->hashtable_num; ++t) keys[(size_t)i * h->hashtable_num + t] = (uint32_t)idx[t];
    }
}

// MultiIndexHashing::insert (:533-598) for entry[i] = feature->mnId with descriptor desc[i], in order.
void orc_mih_insert(orc_mih* h, const uint8_t* desc, const int32_t* entry, int n, uint32_t* hashed, uint8_t* success) {
    std::vector<size_t> idx(h->hashtable_num);
    for (int i = 0; i < n; ++i) {
        bool insert_success = false;
        unsigned d[8];
        std::memcpy(d, desc + 32 * (size_t)i, 32);
        h->hash(d, idx);
        uint32_t flags = hashed ? hashed[i] : 0u;
        for (unsigned t = 0; t < h->hashtable_num; ++t) {
            if (!(flags >> t & 1u)) {                          // feature->mvbHashed[hashtable_id]
                flags |= 1u << t;
                Bucket& b = h->tables[t][idx[t]];
                if (!b.has_latest || b.latest != entry[i]) {   // latest == NULL || latest->mnId != feature->mnId
                    insert_success = true;
                    b.add(entry[i]);
                }
            }
        }
        if (hashed) hashed[i] = flags;
        if (success) success[i] = insert_success;
    }
}

// query(pDesc, candidates, nactive_tables_top, nretrieves) (:699-846): cand[q * stride ..], count[q]; returns the largest count
int orc_mih_query(const orc_mih* h, const uint8_t* qdesc, int m, int nactive, int nretrieve, int32_t* cand, int stride, int32_t* count) {
    std::vector<size_t> idx(h->hashtable_num);
    std::vector<int> entries;
    if (nactive == 0) nactive = (int)h->hashtable_num;         // nactive_tables_top == NULL
    int worst = 0;
    for (int q = 0; q < m; ++q) {
        unsigned d[8];
        std::memcpy(d, qdesc + 32 * (size_t)q, 32);
        h->hash(d, idx);
        int c = 0;
        for (int t = 0; t < nactive; ++t) {
            h->tables[t][idx[t]].back(entries, nretrieve + h->dynamics);
            for (int e : entries) { if (c < stride) cand[(size_t)q * stride + c] = e; ++c; }
        }
        count[q] = c;
        worst = std::max(worst, c);
    }
    return worst;
}

// query(pDesc, candidates, addi_candidates, nfeatures) (:849-950)
int orc_mih_query2(const orc_mih* h, const uint8_t* qdesc, int m, const int32_t* nfeatures, int32_t* cand, int32_t* count,
                   int32_t* addi, int32_t* addi_count, int stride) {
    std::vector<size_t> idx(h->hashtable_num);
    std::vector<int> entries;
    const int keep = kNumCandidateFeatures + h->dynamics;
    int worst = 0;
    for (int q = 0; q < m; ++q) {
        unsigned d[8];
        std::memcpy(d, qdesc + 32 * (size_t)q, 32);
        h->hash(d, idx);
        int c = 0, a = 0;
        for (unsigned k = 0; k < h->hashtable_num; ++k) {
            if (nfeatures[k] > 0) {
                h->tables[k][idx[k]].back(entries, nfeatures[k] + h->dynamics);
                const int s = (int)entries.size();
                if (s <= keep) {
                    for (int e : entries) { if (c < stride) cand[(size_t)q * stride + c] = e; ++c; }
                } else {
                    for (int i = s - keep; i < s; ++i) { if (c < stride) cand[(size_t)q * stride + c] = entries[i]; ++c; }
                    for (int i = 0; i < s - keep; ++i) { if (a < stride) addi[(size_t)q * stride + a] = entries[i]; ++a; }
                }
            }
        }
        count[q] = c; addi_count[q] = a;
        worst = std::max(worst, std::max(c, a));
    }
    return worst;
}

int orc_mih_update_dynamics(orc_mih* h, double e) {            // :957-971
    if (std::abs(e) < kMinTimeDiff) return h->dynamics;
    if (e > 0 && h->dynamics + kNumCandidateFeatures > kMinBucketSize) h->dynamics--;
    if (e < 0 && h->dynamics + kNumCandidateFeatures < kMaxBucketSize) h->dynamics++;
    return h->dynamics;
}

}  // extern "C"
