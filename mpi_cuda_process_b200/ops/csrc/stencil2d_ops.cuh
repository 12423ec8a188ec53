// Per-operator cell arithmetic of the fused 2-D kernel (stencil2d_fused.cu): how a lane's 16-byte vector of cells is
// kept in registers, how one row is updated from its three source rows, the column masks and the residual terms.
// Reference arithmetic: run_mdf (MDF_kernel.cu:10-22) and game_of_life (kernel.cu:10-68); see stencil_ops.cuh.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "stencil_ops.cuh"

namespace mcp {

// ------------------------------------------------------------------------------------------ operators

// Column masks of the scalar-cell operators: one bit per vector component.
template <int VEC_>
struct ComponentMask {
    using Mask = unsigned;
    static constexpr int CPE = 1;  // cells per storage element
    static __device__ __forceinline__ Mask full_mask() { return 0xfu; }
    static __device__ __forceinline__ bool any(Mask m) { return m != 0u; }
    // component v of the lane's vector is cell cx + v
    static __device__ __forceinline__ Mask col_mask(int cx, int xlo, int xhi) {
        Mask m = 0;
#pragma unroll
        for (int v = 0; v < VEC_; ++v) {
            const int gx = cx + v;
            if (gx < xlo || gx > xhi) m |= 1u << v;
        }
        return m;
    }
};

// Adapters for the operators whose per-row state carries the raw vector in `.v`: the row loop talks to every operator
// through sel / to_row / to_vec / acc so that MdfF32 can keep its rows in packed pairs (below).
template <class D, class V_>
struct PlainRows {
    using Out = V_;
    template <class M, class R>
    static __device__ __forceinline__ Out sel(const M& m, const R& keep, const Out& upd) {
        return D::select(m, keep.v, upd);
    }
    static __device__ __forceinline__ auto to_row(const Out& o) { return D::make_row(o); }
    static __device__ __forceinline__ V_ to_vec(const Out& o) { return o; }
    template <class R>
    static __device__ __forceinline__ void acc(const Out& a, const R& b, double& ss, double& mx) {
        D::accum(a, b.v, ss, mx);
    }
};

// Scalar fp32 MDF (FADD/FFMA): kept as the comparison point for the packed operator below (-DMCP_F32_SCALAR).
struct MdfF32Scalar : ComponentMask<4>, PlainRows<MdfF32Scalar, float4> {
    using T = float;
    using V = float4;
    static constexpr int VEC = 4;
    struct Row {
        V v;
    };
    static __device__ __forceinline__ Row make_row(const V& raw) { return Row{raw}; }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        const float wl = __shfl_up_sync(0xffffffffu, c.v.w, 1);
        const float er = __shfl_down_sync(0xffffffffu, c.v.x, 1);
        V o;
        o.x = mdf_cell(c.v.x, n.v.x, s.v.x, wl, c.v.y);
        o.y = mdf_cell(c.v.y, n.v.y, s.v.y, c.v.x, c.v.z);
        o.z = mdf_cell(c.v.z, n.v.z, s.v.z, c.v.y, c.v.w);
        o.w = mdf_cell(c.v.w, n.v.w, s.v.w, c.v.z, er);
        return o;
    }
    static __device__ __forceinline__ V select(unsigned m, const V& keep, const V& upd) {
        V o;
        o.x = (m & 1u) ? keep.x : upd.x;
        o.y = (m & 2u) ? keep.y : upd.y;
        o.z = (m & 4u) ? keep.z : upd.z;
        o.w = (m & 8u) ? keep.w : upd.w;
        return o;
    }
    // residual contribution of 4 cells: differences and their squares in fp32 (the field's own precision),
    // one conversion per vector, accumulation across rows in fp64
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const float d0 = __fsub_rn(a.x, b.x), d1 = __fsub_rn(a.y, b.y);
        const float d2 = __fsub_rn(a.z, b.z), d3 = __fsub_rn(a.w, b.w);
        float q = __fmul_rn(d0, d0);
        q = __fmaf_rn(d1, d1, q);
        q = __fmaf_rn(d2, d2, q);
        q = __fmaf_rn(d3, d3, q);
        ss += (double)q;
        const float m = fmaxf(fmaxf(fabsf(d0), fabsf(d1)), fmaxf(fabsf(d2), fabsf(d3)));
        mx = fmax(mx, (double)m);
    }
};

// fp32 MDF on Blackwell's packed fp32 pipe: add.rn.f32x2 / fma.rn.f32x2 (SASS FADD2 / FFMA2) round each half to
// nearest-even exactly like the scalar instructions, so the result is still the reference's rounding sequence
// (MDF_kernel.cu:20) bit for bit, at half the issue slots.  A lane's four cells (x, y, z, w) live in two 64-bit
// registers p0 = (x, z), p1 = (y, w): then the east operand of p0 IS p1 and the west operand of p1 IS p0, and only the
// two pairs that contain a shuffled neighbour have to be assembled.
struct MdfF32Packed : ComponentMask<4> {
    using T = float;
    using V = float4;
    static constexpr int VEC = 4;
    using u64 = unsigned long long;
    struct Row {
        u64 p0, p1;  // (x, z), (y, w)
    };
    using Out = Row;
    static __device__ __forceinline__ u64 pack(float lo, float hi) {
        u64 d;
        asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
        return d;
    }
    static __device__ __forceinline__ void unpack(u64 v, float& lo, float& hi) {
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    }
    static __device__ __forceinline__ u64 add2(u64 a, u64 b) {
        u64 d;
        asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
        return d;
    }
    static __device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
        u64 d;
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
        return d;
    }
    static __device__ __forceinline__ Row make_row(const V& raw) { return Row{pack(raw.x, raw.z), pack(raw.y, raw.w)}; }
    static __device__ __forceinline__ Row to_row(const Out& o) { return o; }
    static __device__ __forceinline__ V to_vec(const Out& o) {
        V v;
        unpack(o.p0, v.x, v.z);
        unpack(o.p1, v.y, v.w);
        return v;
    }
    // mdf_cell (stencil_ops.cuh) on two cells at once: s = E + W; s += N; s += S; s = fma(C, -4, s); fma(s, 0.25, C)
    static __device__ __forceinline__ u64 cell2(u64 C, u64 N, u64 S, u64 W, u64 E) {
        const u64 m4 = pack(-4.0f, -4.0f), q = pack(0.25f, 0.25f);
        u64 s = add2(E, W);
        s = add2(s, N);
        s = add2(s, S);
        s = fma2(C, m4, s);
        return fma2(s, q, C);
    }
    static __device__ __forceinline__ Out update(const Row& n, const Row& c, const Row& s) {
        float x, y, z, w;
        unpack(c.p0, x, z);
        unpack(c.p1, y, w);
        const float wl = __shfl_up_sync(0xffffffffu, w, 1);
        const float er = __shfl_down_sync(0xffffffffu, x, 1);
        Out o;
        o.p0 = cell2(c.p0, n.p0, s.p0, pack(wl, y), c.p1);  // cells x, z: west (wl, y), east (y, w)
        o.p1 = cell2(c.p1, n.p1, s.p1, c.p0, pack(z, er));  // cells y, w: west (x, z), east (z, er)
        return o;
    }
    static __device__ __forceinline__ Out sel(unsigned m, const Row& keep, const Out& upd) {
        const V k = to_vec(keep), u = to_vec(upd);
        return Row{pack((m & 1u) ? k.x : u.x, (m & 4u) ? k.z : u.z), pack((m & 2u) ? k.y : u.y, (m & 8u) ? k.w : u.w)};
    }
    // residual contribution of 4 cells: differences and their squares in fp32 (the field's own precision),
    // one conversion per vector, accumulation across rows in fp64
    static __device__ __forceinline__ void acc(const Out& a_, const Row& b_, double& ss, double& mx) {
        const V a = to_vec(a_), b = to_vec(b_);
        const float d0 = __fsub_rn(a.x, b.x), d1 = __fsub_rn(a.y, b.y);
        const float d2 = __fsub_rn(a.z, b.z), d3 = __fsub_rn(a.w, b.w);
        float q = __fmul_rn(d0, d0);
        q = __fmaf_rn(d1, d1, q);
        q = __fmaf_rn(d2, d2, q);
        q = __fmaf_rn(d3, d3, q);
        ss += (double)q;
        const float m = fmaxf(fmaxf(fabsf(d0), fabsf(d1)), fmaxf(fabsf(d2), fabsf(d3)));
        mx = fmax(mx, (double)m);
    }
};

#ifdef MCP_F32_SCALAR
using MdfF32 = MdfF32Scalar;
#else
using MdfF32 = MdfF32Packed;
#endif

struct MdfF64 : ComponentMask<2>, PlainRows<MdfF64, double2> {
    using T = double;
    using V = double2;
    static constexpr int VEC = 2;
    struct Row {
        V v;
    };
    static __device__ __forceinline__ Row make_row(const V& raw) { return Row{raw}; }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        const double wl = __shfl_up_sync(0xffffffffu, c.v.y, 1);
        const double er = __shfl_down_sync(0xffffffffu, c.v.x, 1);
        V o;
        o.x = mdf_cell(c.v.x, n.v.x, s.v.x, wl, c.v.y);
        o.y = mdf_cell(c.v.y, n.v.y, s.v.y, c.v.x, er);
        return o;
    }
    static __device__ __forceinline__ V select(unsigned m, const V& keep, const V& upd) {
        V o;
        o.x = (m & 1u) ? keep.x : upd.x;
        o.y = (m & 2u) ? keep.y : upd.y;
        return o;
    }
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const double d0 = a.x - b.x, d1 = a.y - b.y;
        ss = fma(d0, d0, ss); ss = fma(d1, d1, ss);
        mx = fmax(mx, fmax(fabs(d0), fabs(d1)));
    }
};

struct LifeI32 : ComponentMask<4>, PlainRows<LifeI32, int4> {
    using T = int;
    using V = int4;
    static constexpr int VEC = 4;
    struct Row {
        V v;   // cells
        V hs;  // horizontal 3-sums (left + self + right)
    };
    static __device__ __forceinline__ Row make_row(const V& raw) {
        const int wl = __shfl_up_sync(0xffffffffu, raw.w, 1);
        const int er = __shfl_down_sync(0xffffffffu, raw.x, 1);
        Row r;
        r.v = raw;
        const int xy = raw.x + raw.y, zw = raw.z + raw.w;
        r.hs.x = wl + xy;
        r.hs.y = xy + raw.z;
        r.hs.z = raw.y + zw;
        r.hs.w = zw + er;
        return r;
    }
    static __device__ __forceinline__ int rule(int n, int c) { return life_rule(n, c); }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        V o;
        o.x = rule(n.hs.x + s.hs.x + c.hs.x - c.v.x, c.v.x);
        o.y = rule(n.hs.y + s.hs.y + c.hs.y - c.v.y, c.v.y);
        o.z = rule(n.hs.z + s.hs.z + c.hs.z - c.v.z, c.v.z);
        o.w = rule(n.hs.w + s.hs.w + c.hs.w - c.v.w, c.v.w);
        return o;
    }
    static __device__ __forceinline__ V select(unsigned m, const V& keep, const V& upd) {
        V o;
        o.x = (m & 1u) ? keep.x : upd.x;
        o.y = (m & 2u) ? keep.y : upd.y;
        o.z = (m & 4u) ? keep.z : upd.z;
        o.w = (m & 8u) ? keep.w : upd.w;
        return o;
    }
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const int d0 = a.x - b.x, d1 = a.y - b.y, d2 = a.z - b.z, d3 = a.w - b.w;
        const int q = d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
        ss += (double)q;
        mx = fmax(mx, q ? 1.0 : 0.0);
    }
};

// Bit-packed Game of Life: one cell per bit, 32 cells per 32-bit word (cell x = bit x%32 of word x/32), so a lane's
// 16-byte vector holds 128 cells and a warp row 4096.  Same rule as LifeI32 (kernel.cu:61,66), evaluated for 32 cells
// at a time with bit-sliced adders.  The optional fast path of SURVEY section 7.1; the reference stores one int per cell.
struct LifeB1 : PlainRows<LifeB1, uint4> {
    using T = unsigned;
    using V = uint4;
    static constexpr int VEC = 4;   // storage elements per lane
    static constexpr int CPE = 32;  // cells per storage element
    using Mask = uint4;             // one bit per cell: 1 = keep the old value
    struct Row {
        V v;       // cells
        V s0, s1;  // per cell: (s1:s0) = west + self + east (0..3)
    };
    static __device__ __forceinline__ Mask full_mask() {
        return make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    }
    static __device__ __forceinline__ bool any(const Mask& m) { return (m.x | m.y | m.z | m.w) != 0u; }
    // bits of the word holding cells [c0, c0+32) that lie outside [xlo, xhi]
    static __device__ __forceinline__ unsigned word_mask(int c0, int xlo, int xhi) {
        unsigned keep = 0u;  // bits inside [xlo, xhi] -> evolve
        const int lo = max(xlo, c0), hi = min(xhi, c0 + 31);
        if (lo <= hi) {
            const int n = hi - lo + 1;
            const unsigned run = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
            keep = run << (lo - c0);
        }
        return ~keep;
    }
    static __device__ __forceinline__ Mask col_mask(int cx, int xlo, int xhi) {
        // cx = index of the lane's first storage element; element e holds cells 32e .. 32e+31
        return make_uint4(word_mask((cx + 0) * 32, xlo, xhi), word_mask((cx + 1) * 32, xlo, xhi),
                          word_mask((cx + 2) * 32, xlo, xhi), word_mask((cx + 3) * 32, xlo, xhi));
    }
    static __device__ __forceinline__ unsigned maj(unsigned a, unsigned b, unsigned c) { return (a & b) | (c & (a ^ b)); }
    static __device__ __forceinline__ Row make_row(const V& raw) {
        const unsigned prevw = __shfl_up_sync(0xffffffffu, raw.w, 1);    // word to the west of .x
        const unsigned nextx = __shfl_down_sync(0xffffffffu, raw.x, 1);  // word to the east of .w
        Row r;
        r.v = raw;
        const unsigned c[4] = {raw.x, raw.y, raw.z, raw.w};
        const unsigned wprev[4] = {prevw, raw.x, raw.y, raw.z};
        const unsigned enext[4] = {raw.y, raw.z, raw.w, nextx};
        unsigned s0[4], s1[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned W = (c[k] << 1) | (wprev[k] >> 31);  // west neighbour of every cell of the word
            const unsigned E = (c[k] >> 1) | (enext[k] << 31);  // east neighbour
            s0[k] = W ^ c[k] ^ E;
            s1[k] = maj(W, c[k], E);
        }
        r.s0 = make_uint4(s0[0], s0[1], s0[2], s0[3]);
        r.s1 = make_uint4(s1[0], s1[1], s1[2], s1[3]);
        return r;
    }
    static __device__ __forceinline__ unsigned word(unsigned n0, unsigned n1, unsigned c0, unsigned c1, unsigned s0,
                                                    unsigned s1, unsigned centre) {
        // total (0..9, centre included) = l + 2*t1 + 4*t2 + 8*t3
        const unsigned l = n0 ^ c0 ^ s0, cl = maj(n0, c0, s0);
        const unsigned h = n1 ^ c1 ^ s1, ch = maj(n1, c1, s1);
        const unsigned t1 = h ^ cl, c2 = h & cl;
        const unsigned t2 = ch ^ c2, t3 = ch & c2;
        // alive next iff total == 3, or total == 4 and alive now  (= 3 neighbours, or 2 neighbours and alive)
        return ~t3 & ((l & t1 & ~t2) | (~l & ~t1 & t2 & centre));
    }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        return make_uint4(word(n.s0.x, n.s1.x, c.s0.x, c.s1.x, s.s0.x, s.s1.x, c.v.x),
                          word(n.s0.y, n.s1.y, c.s0.y, c.s1.y, s.s0.y, s.s1.y, c.v.y),
                          word(n.s0.z, n.s1.z, c.s0.z, c.s1.z, s.s0.z, s.s1.z, c.v.z),
                          word(n.s0.w, n.s1.w, c.s0.w, c.s1.w, s.s0.w, s.s1.w, c.v.w));
    }
    static __device__ __forceinline__ V select(const Mask& m, const V& keep, const V& upd) {
        return make_uint4((keep.x & m.x) | (upd.x & ~m.x), (keep.y & m.y) | (upd.y & ~m.y),
                          (keep.z & m.z) | (upd.z & ~m.z), (keep.w & m.w) | (upd.w & ~m.w));
    }
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const int q = __popc(a.x ^ b.x) + __popc(a.y ^ b.y) + __popc(a.z ^ b.z) + __popc(a.w ^ b.w);
        ss += (double)q;  // changed cells: (a-b)^2 is 0 or 1
        mx = fmax(mx, q ? 1.0 : 0.0);
    }
};

template <class V>
__device__ __forceinline__ V bitcast_v(const uint4& u) {
    static_assert(sizeof(V) == 16, "vector must be 16 bytes");
    V v;
    memcpy(&v, &u, 16);
    return v;
}
template <class V>
__device__ __forceinline__ uint4 bitcast_u(const V& v) {
    uint4 u;
    memcpy(&u, &v, 16);
    return u;
}

__host__ __device__ constexpr int mod3(int x) { return ((x % 3) + 3) % 3; }

}  // namespace mcp
