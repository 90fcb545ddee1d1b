// Suffix array construction by induced sorting (SA-IS, Nong/Zhang/Chan 2009) over an integer
// alphabet, used by the data-prep tools to build the multi-string BWT the reference consumes
// (the reference delegates this to deps/grlBWT, which needs SDSL and cannot be built here).
#pragma once
#include <cstdint>
#include <vector>

namespace pgi_prep {

template <typename I>
struct SAIS {
    static inline bool tget(const std::vector<uint64_t> &t, I i) { return (t[(uint64_t) i >> 6] >> ((uint64_t) i & 63)) & 1; }
    static inline void tset(std::vector<uint64_t> &t, I i, bool v) {
        if (v) t[(uint64_t) i >> 6] |= 1ULL << ((uint64_t) i & 63); else t[(uint64_t) i >> 6] &= ~(1ULL << ((uint64_t) i & 63));
    }
    template <typename S>
    static void buckets(const S *s, std::vector<I> &bkt, I n, I K, bool end) {
        for (I i = 0; i < K; ++i) bkt[i] = 0;
        for (I i = 0; i < n; ++i) bkt[s[i]]++;
        I sum = 0;
        for (I i = 0; i < K; ++i) { sum += bkt[i]; bkt[i] = end ? sum : sum - bkt[i]; }
    }
    template <typename S>
    static void induce(const std::vector<uint64_t> &t, I *SA, const S *s, std::vector<I> &bkt, I n, I K) {
        buckets(s, bkt, n, K, false);
        for (I i = 0; i < n; ++i) { I j = SA[i] - 1; if (SA[i] > 0 && !tget(t, j)) SA[bkt[s[j]]++] = j; }
        buckets(s, bkt, n, K, true);
        for (I i = n - 1; i >= 0; --i) { I j = SA[i] - 1; if (SA[i] > 0 && tget(t, j)) SA[--bkt[s[j]]] = j; }
    }
    // s[n-1] must be the unique smallest symbol (0); symbols in [0,K)
    template <typename S>
    static void run(const S *s, I *SA, I n, I K) {
        if (n == 1) { SA[0] = 0; return; }
        std::vector<uint64_t> t(((uint64_t) n >> 6) + 1, 0);
        tset(t, n - 1, true); tset(t, n - 2, false);
        for (I i = n - 3; i >= 0; --i) tset(t, i, s[i] < s[i + 1] || (s[i] == s[i + 1] && tget(t, i + 1)));
        auto isLMS = [&](I i) { return i > 0 && tget(t, i) && !tget(t, i - 1); };
        std::vector<I> bkt((size_t) K);
        buckets(s, bkt, n, K, true);
        for (I i = 0; i < n; ++i) SA[i] = -1;
        for (I i = 1; i < n; ++i) if (isLMS(i)) SA[--bkt[s[i]]] = i;
        induce(t, SA, s, bkt, n, K);
        I n1 = 0;
        for (I i = 0; i < n; ++i) if (isLMS(SA[i])) SA[n1++] = SA[i];
        for (I i = n1; i < n; ++i) SA[i] = -1;
        I name = 0, prev = -1;
        for (I i = 0; i < n1; ++i) {
            I pos = SA[i]; bool diff = false;
            for (I d = 0; d < n; ++d) {
                if (prev == -1 || s[pos + d] != s[prev + d] || tget(t, pos + d) != tget(t, prev + d)) { diff = true; break; }
                else if (d > 0 && (isLMS(pos + d) || isLMS(prev + d))) break;
            }
            if (diff) { name++; prev = pos; }
            SA[n1 + pos / 2] = name - 1;
        }
        for (I i = n - 1, j = n - 1; i >= n1; --i) if (SA[i] >= 0) SA[j--] = SA[i];
        I *SA1 = SA, *s1 = SA + n - n1;
        if (name < n1) run<I>(s1, SA1, n1, name);
        else for (I i = 0; i < n1; ++i) SA1[s1[i]] = i;
        buckets(s, bkt, n, K, true);
        for (I i = 1, j = 0; i < n; ++i) if (isLMS(i)) s1[j++] = i;
        for (I i = 0; i < n1; ++i) SA1[i] = s1[SA1[i]];
        for (I i = n1; i < n; ++i) SA[i] = -1;
        for (I i = n1 - 1; i >= 0; --i) { I j = SA[i]; SA[i] = -1; SA[--bkt[s[j]]] = j; }
        induce(t, SA, s, bkt, n, K);
    }
};

} // namespace pgi_prep
