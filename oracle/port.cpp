// TEST INFRASTRUCTURE ONLY. "port" oracle: a plain, single-file CPU restatement of EPIK's placement
// hot path, written from the behaviour of the reference (not a copy of it). Each function cites the
// reference lines it restates (paths relative to /root/reference). It is validated against
// oracle/_ref (the reference's own place.cpp compiled unmodified) by tests/test_oracle.py, and against
// the KAT-1 vectors of SURVEY.md Appendix D.
//
// PARITY: accumulate / threshold default / top-k / LWR are pinned by the reference's code.
// The k-mer codec, DB lookup and load filter restate phylo42/i2l @ 43805ff from its published
// behaviour; that submodule is absent, so those steps are PARITY UNPINNED.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <string_view>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#include <omp.h>

#include "oracle_capi.h"

namespace
{
thread_local std::string g_error;

// ---------------------------------------------------------------- alphabet (i2l, Appendix B) -----
// Per letter: number of resolutions (0 = gap/unknown, 1 = plain, >1 = ambiguous) and their codes.
struct letter_info { uint8_t n; uint8_t codes[20]; };

void set_letter(letter_info* t, char c, std::initializer_list<int> codes)
{
    for (char cc : { c, static_cast<char>(c | 0x20) })
    {
        auto& l = t[static_cast<unsigned char>(cc)];
        l.n = static_cast<uint8_t>(codes.size());
        int i = 0;
        for (int code : codes) { l.codes[i++] = static_cast<uint8_t>(code); }
    }
}

void fill_dna(letter_info* t)
{
    set_letter(t, 'A', {0}); set_letter(t, 'C', {1}); set_letter(t, 'G', {2});
    set_letter(t, 'T', {3}); set_letter(t, 'U', {3});
    set_letter(t, 'R', {0, 2}); set_letter(t, 'Y', {1, 3}); set_letter(t, 'S', {1, 2});
    set_letter(t, 'W', {0, 3}); set_letter(t, 'K', {2, 3}); set_letter(t, 'M', {0, 1});
    set_letter(t, 'B', {1, 2, 3}); set_letter(t, 'D', {0, 2, 3}); set_letter(t, 'H', {0, 1, 3});
    set_letter(t, 'V', {0, 1, 2}); set_letter(t, 'N', {0, 1, 2, 3});
}

void fill_aa(letter_info* t)
{
    const char* order = "RHKDESTNQCGPAILMFWYV";
    for (int i = 0; i < 20; ++i) { set_letter(t, order[i], {i}); }
    set_letter(t, 'U', {9}); set_letter(t, 'O', {14});
    set_letter(t, 'B', {3, 7}); set_letter(t, 'Z', {4, 8}); set_letter(t, 'J', {13, 14});
    set_letter(t, 'X', {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19});
}

// ------------------------------------------------------------------------------- database -----
struct port_db
{
    uint32_t alphabet, bits, k, n_nodes;
    float omega, threshold, log_threshold;
    letter_info letters[256];
    // open addressing: slot -> index into keys (or -1)
    std::vector<int64_t> slots;
    uint64_t mask;
    std::vector<uint32_t> keys;
    std::vector<uint64_t> offsets;
    std::vector<orc_entry> entries;
    std::vector<double> distal, pendant;
    uint64_t keep_at_most;
    double keep_factor;
    uint64_t max_threads;

    uint64_t bucket(uint32_t key) const { return ((key * 0x9E3779B97F4A7C15ull) >> 24) & mask; }

    // i2l::phylo_kmer_db::search, call sites place.cpp:300,311. Returns key index or -1.
    int64_t find(uint32_t key) const
    {
        for (uint64_t i = bucket(key);; i = (i + 1) & mask)
        {
            const int64_t s = slots[i];
            if (s < 0 || keys[static_cast<size_t>(s)] == key) { return s; }
        }
    }
};

// One yielded k-mer key: window start, code, number of keys of its window (1 = unambiguous).
struct kmer_key { uint32_t pos, key, group; };

// i2l::to_kmers<one_ambiguity_policy>, call site place.cpp:294: a window is kept if it has no gap
// and at most one ambiguous letter; an ambiguous window yields one key per resolution.
void enumerate_kmers(const port_db& db, std::string_view seq, std::vector<kmer_key>& out)
{
    out.clear();
    const size_t k = db.k;
    if (seq.size() < k) { return; }
    for (size_t pos = 0; pos + k <= seq.size(); ++pos)
    {
        uint32_t base = 0;
        int n_amb = 0;
        size_t amb_at = 0;
        bool skip = false;
        for (size_t i = 0; i < k && !skip; ++i)
        {
            const auto& l = db.letters[static_cast<unsigned char>(seq[pos + i])];
            base <<= db.bits;
            if (l.n == 0) { skip = true; }
            else if (l.n == 1) { base |= l.codes[0]; }
            else { ++n_amb; amb_at = i; skip = n_amb > 1; }
        }
        if (skip) { continue; }
        if (n_amb == 0)
        {
            out.push_back({ static_cast<uint32_t>(pos), base, 1u });
            continue;
        }
        const auto& l = db.letters[static_cast<unsigned char>(seq[pos + amb_at])];
        const unsigned shift = db.bits * static_cast<unsigned>(k - 1 - amb_at);
        for (unsigned r = 0; r < l.n; ++r)
        {
            out.push_back({ static_cast<uint32_t>(pos), base | (static_cast<uint32_t>(l.codes[r]) << shift), l.n });
        }
    }
}

// ------------------------------------------------------------------------------ placement -----
struct cand
{
    uint32_t branch;
    float score;
    double lwr;
    uint64_t count;
    double distal, pendant;
};

struct scratch
{
    std::vector<float> S, S_amb;
    std::vector<uint64_t> C, C_amb;
    std::vector<uint32_t> touched;
    std::vector<kmer_key> kmers;
};

// place.cpp:320-440 (place_seq): returns one candidate per touched branch, in first-touch order.
std::vector<cand> score_branches(const port_db& db, scratch& w, std::string_view seq)
{
    const uint64_t n_windows = static_cast<uint64_t>(seq.size()) - db.k + 1; // wraps like place.cpp:322

    for (uint32_t b : w.touched)  // sparse reset, place.cpp:335-342
    {
        w.S[b] = 0.0f; w.S_amb[b] = 0.0f; w.C[b] = 0; w.C_amb[b] = 0;
    }
    w.touched.clear();

    enumerate_kmers(db, seq, w.kmers);

    // exact k-mers first, in window order (place.cpp:349-371)
    for (const auto& km : w.kmers)
    {
        if (km.group != 1) { continue; }
        const int64_t idx = db.find(km.key);
        if (idx < 0) { continue; }
        for (uint64_t e = db.offsets[idx]; e < db.offsets[idx + 1]; ++e)
        {
            const uint32_t b = db.entries[e].branch;
            if (w.C[b] == 0) { w.touched.push_back(b); }
            w.C[b] += 1;
            w.S[b] += db.entries[e].score;
        }
    }

    // then every resolved key of the ambiguous windows, each on its own (place.cpp:373-415):
    // a branch first seen by this key gets (10^s + (k - C_amb)·thr)/k added to its LOG score
    // (replicated as is).
    for (const auto& km : w.kmers)
    {
        if (km.group == 1) { continue; }
        const int64_t idx = db.find(km.key);
        if (idx < 0) { continue; }
        std::unordered_set<uint32_t> fresh;
        for (uint64_t e = db.offsets[idx]; e < db.offsets[idx + 1]; ++e)
        {
            const uint32_t b = db.entries[e].branch;
            if (w.C_amb[b] == 0) { fresh.insert(b); }
            w.C_amb[b] += 1;
            w.S_amb[b] += static_cast<float>(std::pow(10, db.entries[e].score)); // double pow, place.cpp:391
        }
        const size_t width = db.k;
        for (const uint32_t b : fresh)
        {
            const float mean = (w.S_amb[b] + static_cast<float>(width - w.C_amb[b]) * db.threshold)
                               / static_cast<float>(width);
            if (w.C[b] == 0) { w.touched.push_back(b); }
            w.C[b] += 1;
            w.S[b] += mean;
        }
    }

    // omega-threshold default for the windows absent on a branch (place.cpp:417-422)
    std::vector<cand> out;
    out.reserve(w.touched.size());
    for (const uint32_t b : w.touched)
    {
        w.S[b] += static_cast<float>(n_windows - w.C[b]) * db.log_threshold;
        w.S[b] /= static_cast<float>(db.k);
        out.push_back({ b, w.S[b], 0.0, w.C[b], db.distal[b], db.pendant[b] }); // place.cpp:424-438
    }
    return out;
}

// place.cpp:164-184 (sum_scores): float factors, double powers, over ALL touched branches.
double likelihood_mass(const port_db& db, const std::vector<cand>& cands, size_t seq_len)
{
    const float n_branches = static_cast<float>(db.n_nodes);
    const float n_placed = static_cast<float>(cands.size());
    const float n_windows = static_cast<float>(static_cast<uint64_t>(seq_len) - db.k + 1);
    const float kf = static_cast<float>(db.k);
    const double untouched = (n_branches - n_placed) * std::pow(10.0, static_cast<double>(n_windows * db.log_threshold / kf));
    double touched = 0.0;
    for (const auto& c : cands) { touched += std::pow(10.0, static_cast<double>(c.score)); }
    return untouched + touched;
}

bool better(const cand& a, const cand& b) { return a.score > b.score; }

// place.cpp:230-268: the per-sequence body of placer::place.
std::vector<cand> place_one(const port_db& db, scratch& w, std::string_view seq)
{
    std::vector<cand> cands = score_branches(db, w, seq);
    const double mass = likelihood_mass(db, cands, seq.size());
    const uint64_t n_windows = static_cast<uint64_t>(seq.size()) - db.k + 1;

    // place.cpp:134-159 (select_best_placements), incl. the zero-hit fallback :141-152
    size_t keep = std::min<size_t>(db.keep_at_most, cands.size());
    if (keep == 0)
    {
        keep = db.keep_at_most;
        const float s = db.log_threshold * static_cast<float>(n_windows) / static_cast<float>(db.k);
        for (size_t i = 0; i < keep; ++i) { cands.push_back({ static_cast<uint32_t>(i), s, 0.0, 0, 0.0, 0.0 }); }
    }
    std::partial_sort(cands.begin(), cands.begin() + static_cast<long>(keep), cands.end(), better);
    cands.resize(keep);

    // place.cpp:241-264 (LWR with its two zero guards)
    double keep_factor = db.keep_factor;
    for (auto& c : cands)
    {
        if (mass == 0) { c.lwr = 0.0; keep_factor = 0.0; continue; }
        const double p = std::pow(10.0, static_cast<double>(c.score));
        c.lwr = (p == 0.0) ? 0.0 : p / mass;
    }

    // place.cpp:188-199 (filter_by_ratio)
    const double cut = (cands.empty() ? 0.0 : cands[0].lwr) * keep_factor;
    std::vector<cand> kept;
    for (const auto& c : cands) { if (c.lwr >= cut) { kept.push_back(c); } }
    return kept;
}

struct port_handle
{
    port_db db;
    std::vector<scratch> scratch_per_thread;
};
} // namespace

extern "C" int orc_alphabet(void) { return 2; }
extern "C" const char* orc_kind(void) { return "port"; }
extern "C" const char* orc_last_error(void) { return g_error.c_str(); }

extern "C" void* orc_create(const orc_db_desc* d, uint64_t keep_at_most, double keep_factor, uint64_t max_threads)
{
    try
    {
        auto* h = new port_handle();
        auto& db = h->db;
        db.alphabet = d->alphabet;
        db.bits = d->alphabet ? 5 : 2;
        db.k = d->kmer_size;
        db.n_nodes = d->n_nodes;
        db.omega = d->omega;
        // i2l::score_threshold (place.cpp:87) = (omega/|alphabet|)^k in float; place.cpp:88 log10 in float
        db.threshold = std::pow(d->omega / static_cast<float>(d->alphabet ? 20 : 4), static_cast<float>(d->kmer_size));
        db.log_threshold = std::log10(db.threshold);
        std::memset(db.letters, 0, sizeof(db.letters));
        if (d->alphabet) { fill_aa(db.letters); } else { fill_dna(db.letters); }
        db.keep_at_most = keep_at_most;
        db.keep_factor = keep_factor;
        db.max_threads = std::max<uint64_t>(max_threads, 1);

        db.keys.assign(d->keys, d->keys + d->n_keys);
        db.offsets.assign(d->offsets, d->offsets + d->n_keys + 1);
        db.entries.assign(d->entries, d->entries + d->offsets[d->n_keys]);
        uint64_t cap = 16;
        while (cap < 2 * d->n_keys + 2) { cap <<= 1; }
        db.slots.assign(cap, -1);
        db.mask = cap - 1;
        for (uint64_t i = 0; i < d->n_keys; ++i)
        {
            uint64_t s = db.bucket(db.keys[i]);
            while (db.slots[s] >= 0)
            {
                if (db.keys[static_cast<size_t>(db.slots[s])] == db.keys[i]) { throw std::runtime_error("duplicate key"); }
                s = (s + 1) & db.mask;
            }
            db.slots[s] = static_cast<int64_t>(i);
        }

        // place.cpp:98-125: distal = branch_length / 2; pendant = mean subtree branch length + distal
        for (uint32_t i = 0; i < d->n_nodes; ++i)
        {
            const double distal = d->branch_length[i] / 2;
            double mean = 0.0;
            if (d->subtree_num_nodes[i] > 1) { mean = d->subtree_total_length[i] / static_cast<double>(d->subtree_num_nodes[i]); }
            db.distal.push_back(distal);
            db.pendant.push_back(mean + distal);
        }

        h->scratch_per_thread.resize(db.max_threads);
        for (auto& w : h->scratch_per_thread)
        {
            w.S.assign(d->n_nodes, 0.0f); w.S_amb.assign(d->n_nodes, 0.0f);
            w.C.assign(d->n_nodes, 0); w.C_amb.assign(d->n_nodes, 0);
        }
        return h;
    }
    catch (const std::exception& e)
    {
        g_error = e.what();
        return nullptr;
    }
}

extern "C" void orc_destroy(void* handle) { delete static_cast<port_handle*>(handle); }

extern "C" void orc_thresholds(void* handle, float* threshold, float* log10_threshold)
{
    const auto* h = static_cast<port_handle*>(handle);
    *threshold = h->db.threshold;
    *log10_threshold = h->db.log_threshold;
}

extern "C" int orc_place(void* handle, const char* seq, const uint64_t* seq_off, uint32_t n,
                         uint32_t num_threads, uint32_t batch_size,
                         uint32_t* n_place, uint32_t* branch, float* score, double* lwr, uint64_t* count,
                         double* distal, double* pendant, uint32_t* group_rep, double* seconds)
{
    auto* h = static_cast<port_handle*>(handle);
    const auto& db = h->db;
    if (num_threads == 0 || num_threads > db.max_threads) { g_error = "bad num_threads"; return 1; }
    if (batch_size == 0) { batch_size = n ? n : 1; }
    const uint64_t K = db.keep_at_most;
    double total = 0.0;
    for (uint32_t first = 0; first < n; first += batch_size)
    {
        const uint32_t last = static_cast<uint32_t>(std::min<uint64_t>(n, static_cast<uint64_t>(first) + batch_size));
        const auto t0 = std::chrono::steady_clock::now();

        // place.cpp:73-81,207-212: identical sequences of a batch are placed once
        std::unordered_map<std::string_view, std::vector<uint32_t>> groups;
        for (uint32_t i = first; i < last; ++i)
        {
            groups[std::string_view(seq + seq_off[i], seq_off[i + 1] - seq_off[i])].push_back(i);
        }
        std::vector<const std::pair<const std::string_view, std::vector<uint32_t>>*> uniq;
        uniq.reserve(groups.size());
        for (const auto& g : groups) { uniq.push_back(&g); }
        std::vector<std::vector<cand>> placed(uniq.size());

        // place.cpp:218-230: dynamic OpenMP loop over the unique sequences
#pragma omp parallel for schedule(dynamic) num_threads(num_threads)
        for (size_t u = 0; u < uniq.size(); ++u)
        {
            placed[u] = place_one(db, h->scratch_per_thread[static_cast<size_t>(omp_get_thread_num())], uniq[u]->first);
        }
        const auto t1 = std::chrono::steady_clock::now();
        total += std::chrono::duration<double>(t1 - t0).count();

        for (size_t u = 0; u < uniq.size(); ++u)
        {
            for (const uint32_t i : uniq[u]->second)
            {
                n_place[i] = static_cast<uint32_t>(placed[u].size());
                if (group_rep) { group_rep[i] = uniq[u]->second.front(); }
                uint64_t j = static_cast<uint64_t>(i) * K;
                for (const auto& c : placed[u])
                {
                    if (branch) { branch[j] = c.branch; }
                    if (score) { score[j] = c.score; }
                    if (lwr) { lwr[j] = c.lwr; }
                    if (count) { count[j] = c.count; }
                    if (distal) { distal[j] = c.distal; }
                    if (pendant) { pendant[j] = c.pendant; }
                    ++j;
                }
            }
        }
    }
    if (seconds) { *seconds = total; }
    return 0;
}

extern "C" uint64_t orc_kmers(void* handle, const char* seq, uint64_t len, uint64_t cap,
                              uint32_t* pos, uint32_t* key, uint32_t* group, int64_t* hit_len)
{
    const auto* h = static_cast<port_handle*>(handle);
    std::vector<kmer_key> kmers;
    enumerate_kmers(h->db, std::string_view(seq, len), kmers);
    uint64_t n = 0;
    for (const auto& km : kmers)
    {
        if (n < cap)
        {
            pos[n] = km.pos; key[n] = km.key; group[n] = km.group;
            const int64_t idx = h->db.find(km.key);
            hit_len[n] = idx < 0 ? -1 : static_cast<int64_t>(h->db.offsets[idx + 1] - h->db.offsets[idx]);
        }
        ++n;
    }
    return n;
}
