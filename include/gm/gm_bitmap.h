// Byte-addressed bitmap helpers (reference gm_bitmap.h / gm_bitmap.cc:3-45): bit n lives in byte n/8,
// position n%8, LSB first.  Host side only; the device BFS keeps the same predicate in 32-bit words.
#ifndef GM_BITMAP_H_
#define GM_BITMAP_H_
#include <stdint.h>
#include "gm_atomic_wrapper.h"
unsigned _gm_get_bit(unsigned char* Bit, int n);
void _gm_set_bit(unsigned char* BitMap, int n);
bool _gm_set_bit_atomic(unsigned char* BitMap, int n);      // true if this call set the bit
void _gm_clear_bit(unsigned char* BitMap, int n);
bool _gm_clear_bit_atomic(unsigned char* BitMap, int n);    // true if this call cleared the bit
#endif
