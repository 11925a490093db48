// tests/cpp/test_cointable.cpp — CPU-only: bcp::CoinTable (include/compat/bcp_chain.h: linear probing with backward-shift
// deletion, growth, inline and long scripts) against std::map under a long random mix of insert / find / erase /
// overwrite, with a tiny initial capacity so that probe runs wrap around the end of the table and growth happens
// many times; and bcp::WorkerTeam (every worker runs every job exactly once).
#include <atomic>
#include <cstdio>
#include <map>
#include <random>
#include <string>

#include "compat/bcp_chain.h"

using namespace bcp;

static OutPoint key_of(uint64_t id, uint32_t vout, bool clustered) {
    OutPoint k;
    memset(k.b, 0, 36);
    // clustered keys share their hash's low bits: long probe runs, the hard case for backward-shift deletion
    uint64_t h = clustered ? (id << 20) | 5 : id * 0x9E3779B97F4A7C15ull;
    memcpy(k.b, &h, 8);
    memcpy(k.b + 8, &id, 8);
    memcpy(k.b + 32, &vout, 4);
    return k;
}

int main() {
    std::mt19937_64 rng(99);
    for (int mode = 0; mode < 2; mode++) {
        CoinTable t(8);
        std::map<std::string, std::pair<std::string, int64_t>> ref;
        const uint64_t universe = mode ? 600 : 5000;
        for (int step = 0; step < 400000; step++) {
            const uint64_t id = rng() % universe;
            const uint32_t vout = (uint32_t)(rng() % 3);
            const OutPoint k = key_of(id, vout, mode == 1);
            const std::string ks((const char *)k.b, 36);
            const int op = (int)(rng() % 10);
            if (op < 4) { // insert / overwrite, sometimes with a script longer than the inline space
                const size_t len = (rng() % 20 == 0) ? 65 + rng() % 200 : rng() % 65;
                std::string spk(len, '\0');
                for (auto &c : spk) c = (char)rng();
                const int64_t amount = (int64_t)(rng() >> 8);
                t.insert(k, (const uint8_t *)spk.data(), len, amount);
                ref[ks] = {spk, amount};
            } else if (op < 8) { // find
                CoinSlot *c = t.find(k);
                auto it = ref.find(ks);
                if ((c != nullptr) != (it != ref.end())) { printf("find mismatch at step %d\n", step); return 1; }
                if (c) {
                    if (c->amount != it->second.second || c->spk_len != it->second.first.size() ||
                        memcmp(CoinTable::script(*c), it->second.first.data(), c->spk_len) != 0) {
                        printf("content mismatch at step %d\n", step);
                        return 1;
                    }
                }
            } else { // erase
                CoinSlot *c = t.find(k);
                auto it = ref.find(ks);
                if ((c != nullptr) != (it != ref.end())) { printf("erase/find mismatch at step %d\n", step); return 1; }
                if (c) { t.erase(c); ref.erase(it); }
            }
            if (t.size() != ref.size()) { printf("size mismatch at step %d: %zu vs %zu\n", step, t.size(), ref.size()); return 1; }
        }
        // everything the reference map holds is still reachable, with its content
        for (const auto &kv : ref) {
            OutPoint k;
            memcpy(k.b, kv.first.data(), 36);
            CoinSlot *c = t.find(k);
            if (!c || c->amount != kv.second.second || c->spk_len != kv.second.first.size() ||
                memcmp(CoinTable::script(*c), kv.second.first.data(), c->spk_len) != 0) {
                printf("final sweep mismatch (mode %d)\n", mode);
                return 1;
            }
        }
        printf("mode %d: %zu coins left after 400000 operations\n", mode, t.size());
    }
    { // WorkerTeam: each job reaches each worker once; many short jobs in a row
        WorkerTeam team(8);
        std::atomic<long> sum{0};
        std::vector<int> seen(8);
        for (int job = 0; job < 2000; job++) {
            std::fill(seen.begin(), seen.end(), 0);
            team.Run([&](int w) { seen[(size_t)w]++; sum += w + 1; });
            for (int w = 0; w < 8; w++)
                if (seen[(size_t)w] != 1) { printf("worker %d ran %d times in job %d\n", w, seen[(size_t)w], job); return 1; }
        }
        if (sum != 2000L * 36) { printf("team sum %ld\n", (long)sum); return 1; }
    }
    printf("cointable ok\n");
    return 0;
}
