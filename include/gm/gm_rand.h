// xorshift generators with the reference's interface (gm_rand.h, gm_rand.cc:10-23).  gm_rand32 keeps the
// reference's second step as a plain shift (`next = next >> 17`, gm_rand.cc:20) so seeded sequences match.
#ifndef GM_RAND_H_
#define GM_RAND_H_
#include <stdint.h>
class gm_rand64 {
  public:
    gm_rand64();
    gm_rand64(int64_t seed) : next(seed) {}
    int64_t rand();
  private:
    int64_t next;
};
class gm_rand32 {
  public:
    gm_rand32() : next((int32_t)2463534242u) {}
    gm_rand32(int32_t seed) : next(seed) {}
    int32_t rand();
  private:
    int32_t next;
};
class gm_rand {
  public:
    gm_rand() {}
    gm_rand(long seed) : rng((int32_t)seed) {}
    int32_t rand() { return rng.rand(); }
  private:
    gm_rand32 rng;
};
#endif
