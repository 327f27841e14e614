// quest-b200: Mersenne Twister MT19937 (Matsumoto & Nishimura 1998, 2002 initialisation), written from the
// published recurrence.  Exports the same C symbols as the reference's QuEST/src/mt19937ar.c
// (init_genrand :61, init_by_array :80, genrand_int32 :106, genrand_real1 :150, ...), because
//   - measurement outcomes must be bit-identical to the reference under the same seedQuEST() key, and
//   - the reference's own known-answer test (tests/essential/state_vector/seedQuEST.test) calls
//     genrand_real1() through the shared library.
// State is process-global, as in the reference (mt19937ar.c:57-58): the API is single-threaded by design.
#include <cstdint>

#include "quest_internal.h"

namespace {

constexpr int kN = 624;          // state words
constexpr int kM = 397;          // middle word offset
constexpr uint32_t kMatrixA = 0x9908b0dfu;
constexpr uint32_t kUpper = 0x80000000u;
constexpr uint32_t kLower = 0x7fffffffu;

struct Twister {
    uint32_t word[kN];
    int next = kN + 1;  // kN+1 flags "never seeded"

    void seed(uint32_t s) {
        word[0] = s;
        for (int i = 1; i < kN; ++i)
            word[i] = 1812433253u * (word[i - 1] ^ (word[i - 1] >> 30)) + static_cast<uint32_t>(i);
        next = kN;
    }

    void seed_by_key(const unsigned long* key, int len) {
        seed(19650218u);
        int i = 1, j = 0;
        for (int k = (kN > len ? kN : len); k > 0; --k) {
            word[i] = (word[i] ^ ((word[i - 1] ^ (word[i - 1] >> 30)) * 1664525u)) +
                      static_cast<uint32_t>(key[j]) + static_cast<uint32_t>(j);
            if (++i >= kN) { word[0] = word[kN - 1]; i = 1; }
            if (++j >= len) j = 0;
        }
        for (int k = kN - 1; k > 0; --k) {
            word[i] = (word[i] ^ ((word[i - 1] ^ (word[i - 1] >> 30)) * 1566083941u)) - static_cast<uint32_t>(i);
            if (++i >= kN) { word[0] = word[kN - 1]; i = 1; }
        }
        word[0] = kUpper;  // guarantees a non-zero initial state
    }

    static uint32_t twist(uint32_t hi, uint32_t lo, uint32_t far) {
        uint32_t y = (hi & kUpper) | (lo & kLower);
        return far ^ (y >> 1) ^ ((y & 1u) ? kMatrixA : 0u);
    }

    void refill() {
        if (next == kN + 1) seed(5489u);  // default seed of the published generator
        int k = 0;
        for (; k < kN - kM; ++k) word[k] = twist(word[k], word[k + 1], word[k + kM]);
        for (; k < kN - 1; ++k) word[k] = twist(word[k], word[k + 1], word[k + kM - kN]);
        word[kN - 1] = twist(word[kN - 1], word[0], word[kM - 1]);
        next = 0;
    }

    uint32_t draw() {
        if (next >= kN) refill();
        uint32_t y = word[next++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        return y;
    }
};

Twister g_mt;

}  // namespace

extern "C" {

void init_genrand(unsigned long s) { g_mt.seed(static_cast<uint32_t>(s)); }
void init_by_array(unsigned long init_key[], int key_length) { g_mt.seed_by_key(init_key, key_length); }
unsigned long genrand_int32(void) { return g_mt.draw(); }
long genrand_int31(void) { return static_cast<long>(g_mt.draw() >> 1); }
double genrand_real1(void) { return g_mt.draw() * (1.0 / 4294967295.0); }          // [0,1]
double genrand_real2(void) { return g_mt.draw() * (1.0 / 4294967296.0); }          // [0,1)
double genrand_real3(void) { return (g_mt.draw() + 0.5) * (1.0 / 4294967296.0); }  // (0,1)
double genrand_res53(void) {
    uint32_t a = g_mt.draw() >> 5, b = g_mt.draw() >> 6;
    return (a * 67108864.0 + b) * (1.0 / 9007199254740992.0);
}

}  // extern "C"
