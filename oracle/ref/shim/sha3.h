/* TEST INFRASTRUCTURE ONLY (oracle/): stand-in for depends/SHA3IUF (github.com/harryr/SHA3IUF @ 3f8de13, absent from the
 * reference tree - .SUBMODULES.json "status": "missing"), exposing the three entry points src/gadgets/mimc.hpp:288-298 uses.
 * Keccak-f[1600] sponge, rate 136, written from the published specification (FIPS 202 / the Keccak reference).
 * The padding byte is a run-time switch because the pinned fork's default is not visible here: 0x01 = original Keccak
 * (Ethereum's keccak256), 0x06 = SHA3-256.  circuit_dump picks the one that reproduces the reference's own Merkle-root
 * known answer (src/test/test_merkle_tree.cpp:56-62). */
#pragma once
#include <stdint.h>
#include <string.h>
#include <stddef.h>

typedef struct { uint64_t s[25]; uint8_t buf[136]; unsigned pos; uint8_t out[32]; } sha3_context;

static unsigned char& b200_shim_sha3_pad() { static unsigned char pad = 0x01; return pad; }

static inline uint64_t b200_shim_rotl(uint64_t x, int n) { return n ? (x << n) | (x >> (64 - n)) : x; }

static inline void b200_shim_keccakf(uint64_t s[25]) {
    static const uint64_t RC[24] = {
        0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL, 0x000000000000808bULL,
        0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL, 0x000000000000008aULL, 0x0000000000000088ULL,
        0x0000000080008009ULL, 0x000000008000000aULL, 0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL,
        0x8000000000008003ULL, 0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
        0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
    static const int ROT[25] = {0, 1, 62, 28, 27, 36, 44, 6, 55, 20, 3, 10, 43, 25, 39, 41, 45, 15, 21, 8, 18, 2, 61, 56, 14};
    for (int round = 0; round < 24; ++round) {
        uint64_t c[5], d[5], b[25];
        for (int x = 0; x < 5; ++x) c[x] = s[x] ^ s[x + 5] ^ s[x + 10] ^ s[x + 15] ^ s[x + 20];
        for (int x = 0; x < 5; ++x) d[x] = c[(x + 4) % 5] ^ b200_shim_rotl(c[(x + 1) % 5], 1);
        for (int i = 0; i < 25; ++i) s[i] ^= d[i % 5];
        for (int x = 0; x < 5; ++x)
            for (int y = 0; y < 5; ++y) b[y + 5 * ((2 * x + 3 * y) % 5)] = b200_shim_rotl(s[x + 5 * y], ROT[x + 5 * y]);
        for (int y = 0; y < 5; ++y)
            for (int x = 0; x < 5; ++x) s[x + 5 * y] = b[x + 5 * y] ^ (~b[(x + 1) % 5 + 5 * y] & b[(x + 2) % 5 + 5 * y]);
        s[0] ^= RC[round];
    }
}

static inline void b200_shim_absorb(sha3_context* c) {
    for (int i = 0; i < 17; ++i) {
        uint64_t w = 0;
        for (int j = 0; j < 8; ++j) w |= (uint64_t)c->buf[8 * i + j] << (8 * j);
        c->s[i] ^= w;
    }
    b200_shim_keccakf(c->s);
    c->pos = 0;
}

static inline void sha3_Init256(sha3_context* c) { memset(c, 0, sizeof *c); }

static inline void sha3_Update(sha3_context* c, const void* data, size_t len) {
    const uint8_t* p = (const uint8_t*)data;
    for (size_t i = 0; i < len; ++i) {
        c->buf[c->pos++] = p[i];
        if (c->pos == 136) b200_shim_absorb(c);
    }
}

static inline const void* sha3_Finalize(sha3_context* c) {
    memset(c->buf + c->pos, 0, 136 - c->pos);
    c->buf[c->pos] ^= b200_shim_sha3_pad();
    c->buf[135] ^= 0x80;
    b200_shim_absorb(c);
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 8; ++j) c->out[8 * i + j] = (uint8_t)(c->s[i] >> (8 * j));
    return c->out;
}
