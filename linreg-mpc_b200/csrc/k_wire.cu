// SecureMultiplication__Msg on the wire (src/phase1.c:100-145, src/protobuf/*.proto): packed
// repeated varints produced and consumed on the GPU, so that the masks and party messages
// never exist as host word arrays.  Both directions are streaming, HBM-bound passes:
//   pack   : k_pack_len (bytes per 2048-word tile) -> k_pack_scan (tile offsets, header,
//            trailer, size) -> k_pack_write (varints assembled in registers, staged in shared
//            memory at the tile's global 16-byte phase, written with aligned 16-byte stores)
//   unpack : k_unpack_count (terminator bytes per 8 KB tile) -> k_unpack_scan (tile offsets,
//            header/trailer checks, value) -> k_unpack_write (decode at every varint start)
// A varint start is the byte after a byte with a clear MSB, so element k of the vector is
// the varint that has exactly k terminator bytes before it in the payload: a prefix count,
// no sequential parse.
#include "p1_common.cuh"

#define WR_THREADS 256
#define PK_WPT 8                          // words per thread
#define PK_TILE (WR_THREADS * PK_WPT)     // words per tile
#define UP_BPT 32                         // bytes per thread
#define UP_TILE (WR_THREADS * UP_BPT)     // bytes per tile


// ------------------------------------------------------------------ device helpers

__device__ __forceinline__ uint32_t vi_len64(uint64_t v) {
	// bytes of the base-128 varint of v: max(1, ceil(bits / 7))
	const uint32_t bits = 64 - __clzll((long long) (v | 1));
	return (bits + 6) / 7;
}

// 28 bits -> four bytes of 7 bits each (two mask-and-shift steps on a 32-bit word) and back
__device__ __forceinline__ uint32_t vi_spread28(uint32_t t) {
	const uint32_t x = (t & 0x00003fffu) | ((t << 2) & 0x3fff0000u);
	return (x & 0x007f007fu) | ((x << 1) & 0x7f007f00u);
}
__device__ __forceinline__ uint32_t vi_gather28(uint32_t y) {
	y = (y & 0x007f007fu) | ((y & 0x7f007f00u) >> 1);
	return (y & 0x00003fffu) | ((y & 0x3fff0000u) >> 2);
}
// 0x80 in the lowest min(c, 4) bytes
__device__ __forceinline__ uint32_t vi_cont4(uint32_t c) {
	return c ? 0x80808080u >> (8 * (4 - min(c, 4u))) : 0u;
}
// the varint of v (L bytes) as three little-endian words: bytes 0-3, 4-7, 8-9; zero beyond byte L-1
__device__ __forceinline__ void vi_encode(uint64_t v, uint32_t L, uint32_t w[3]) {
	const uint32_t lo = (uint32_t) v, hi = (uint32_t) (v >> 32), c = L - 1, t2 = hi >> 24;
	w[0] = vi_spread28(lo & 0x0fffffffu) | vi_cont4(c);
	w[1] = vi_spread28(__funnelshift_r(lo, hi, 28) & 0x0fffffffu) | vi_cont4(c > 4 ? c - 4 : 0);
	w[2] = (t2 & 0x7fu) | ((t2 & 0x80u) << 1) | (c > 8 ? 0x80u : 0u);
}

// the thread's PK_WPT consecutive words of message p, zero past n (vector loads: rows are
// 16-byte aligned, ld * sizeof(W) is a multiple of 16)
template <typename W>
__device__ __forceinline__ void pk_load(const W *__restrict__ row, uint64_t first, uint64_t n, uint64_t v[PK_WPT]) {
	constexpr int PER = 16 / sizeof(W);
#pragma unroll
	for(int c = 0; c < PK_WPT / PER; c++) {
		const uint64_t k = first + (uint64_t) c * PER;
		if(k + PER <= n) {
			const uint4 q = __ldcs(reinterpret_cast<const uint4 *>(row + k));
			if(sizeof(W) == 8) {
				v[c * PER] = ((uint64_t) q.y << 32) | q.x;
				v[c * PER + 1] = ((uint64_t) q.w << 32) | q.z;
			} else {
				v[c * PER] = q.x; v[c * PER + 1] = q.y; v[c * PER + 2] = q.z; v[c * PER + 3] = q.w;
			}
		} else {
#pragma unroll
			for(int e = 0; e < PER; e++) v[c * PER + e] = k + e < n ? (uint64_t) row[k + e] : 0;
		}
	}
}

// ------------------------------------------------------------------ pack

struct PackParams {
	const void *vec;
	uint64_t ld, n;
	uint32_t tiles;
	const uint64_t *values;
	uint8_t *out;
	uint64_t stride;
	uint64_t *sizes;
	uint32_t *tile_sum;   // [count][tiles] payload bytes of a tile
	uint64_t *tile_off;   // [count][tiles] byte offset of the tile's first varint inside the message
};

template <typename W>
__global__ void __launch_bounds__(WR_THREADS) k_pack_len(const PackParams P) {
	__shared__ uint32_t s_warp[WR_THREADS / 32 + 1];
	const uint32_t p = blockIdx.y, tile = blockIdx.x;
	const W *row = (const W *) P.vec + (uint64_t) p * P.ld;
	const uint64_t first = (uint64_t) tile * PK_TILE + threadIdx.x * PK_WPT;
	uint64_t v[PK_WPT];
	pk_load<W>(row, first, P.n, v);
	uint32_t len = 0;
#pragma unroll
	for(int e = 0; e < PK_WPT; e++) len += first + e < P.n ? vi_len64(v[e]) : 0;
#pragma unroll
	for(int o = 16; o; o >>= 1) len += __shfl_xor_sync(0xffffffffu, len, o);
	if((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = len;
	__syncthreads();
	if(threadIdx.x == 0) {
		uint32_t t = 0;
		for(int i = 0; i < WR_THREADS / 32; i++) t += s_warp[i];
		P.tile_sum[(uint64_t) p * P.tiles + tile] = t;
	}
}

__device__ __forceinline__ uint32_t vi_put_global(uint8_t *out, uint64_t v) {
	uint32_t s = 0;
	for(; v >= 0x80; v >>= 7) out[s++] = (uint8_t) (v | 0x80);
	out[s++] = (uint8_t) v;
	return s;
}

// one CTA per message: tile offsets, header [0x0A][len], trailer [0x10][value], size
__global__ void __launch_bounds__(WR_THREADS) k_pack_scan(const PackParams P, int ring_bits) {
	__shared__ uint64_t s_part[WR_THREADS];
	__shared__ uint64_t s_total;
	const uint32_t p = blockIdx.x, t = threadIdx.x;
	const uint32_t per = (P.tiles + WR_THREADS - 1) / WR_THREADS;
	const uint32_t lo = min(t * per, P.tiles), hi = min(lo + per, P.tiles);
	const uint32_t *ts = P.tile_sum + (uint64_t) p * P.tiles;
	uint64_t *to = P.tile_off + (uint64_t) p * P.tiles;
	uint64_t sum = 0;
	for(uint32_t i = lo; i < hi; i++) sum += ts[i];
	s_part[t] = sum;
	__syncthreads();
	if(t == 0) {
		uint64_t run = 0;
		for(int i = 0; i < WR_THREADS; i++) { const uint64_t c = s_part[i]; s_part[i] = run; run += c; }
		s_total = run;
	}
	__syncthreads();
	const uint64_t payload = s_total;
	uint32_t hdr = 0;
	if(P.n) { hdr = 1; for(uint64_t c = payload; ; c >>= 7) { hdr++; if(c < 0x80) break; } }
	uint64_t run = s_part[t] + hdr;
	for(uint32_t i = lo; i < hi; i++) { to[i] = run; run += ts[i]; }
	if(t == 0) {
		uint8_t *out = P.out + (uint64_t) p * P.stride;
		uint64_t o = 0;
		if(P.n) {
			out[o++] = 0x0A;
			o += vi_put_global(out + o, payload);
			o += payload;
		}
		out[o++] = 0x10;
		uint64_t val = P.values[p];
		if(ring_bits == 32) val &= 0xffffffffull;
		o += vi_put_global(out + o, val);
		P.sizes[p] = o;
	}
}

// Varints of the tile, assembled per thread in a 64-bit shift register and emitted as
// aligned 32-bit shared-memory stores (byte stores only for a thread's first / last partial
// word); the staged tile sits at the same 16-byte phase as its place in the message, so the
// copy-out is aligned 16-byte global stores except for the two boundary chunks.
template <typename W>
__global__ void __launch_bounds__(WR_THREADS) k_pack_write(const PackParams P) {
	constexpr int MAXLEN = sizeof(W) == 8 ? 10 : 5;
	__shared__ __align__(16) uint32_t stage[(PK_TILE * MAXLEN + 32) / 4];
	__shared__ uint32_t s_warp[WR_THREADS / 32 + 1];
	const uint32_t p = blockIdx.y, tile = blockIdx.x;
	const W *row = (const W *) P.vec + (uint64_t) p * P.ld;
	const uint64_t first = (uint64_t) tile * PK_TILE + threadIdx.x * PK_WPT;
	uint64_t v[PK_WPT];
	pk_load<W>(row, first, P.n, v);
	uint32_t len[PK_WPT], tot = 0;
#pragma unroll
	for(int e = 0; e < PK_WPT; e++) {
		len[e] = first + e < P.n ? vi_len64(v[e]) : 0;
		tot += len[e];
	}
	uint32_t tile_bytes;
	const uint32_t excl = p1_block_excl_scan256(tot, s_warp, &tile_bytes);
	const uint64_t goff = P.tile_off[(uint64_t) p * P.tiles + tile];
	const uint32_t gshift = (uint32_t) goff & 15u;
	uint8_t *sb = reinterpret_cast<uint8_t *>(stage);

	// the thread's varints are appended to a word-aligned stream: `carry` holds the `fill` (0..3)
	// bytes not yet written, every varint contributes up to three words through funnel shifts
	const uint32_t pos = gshift + excl, head = pos & 3u;
	uint32_t wi = pos >> 2, fill = head, carry = 0;
	bool partial = head != 0;      // the first word is shared with the previous thread: byte stores
	auto emit = [&](uint32_t word) {
		if(partial) {
			for(uint32_t b = head; b < 4; b++) sb[wi * 4 + b] = (uint8_t) (word >> (8 * b));
			partial = false;
		} else stage[wi] = word;
		wi++;
	};
#pragma unroll
	for(int e = 0; e < PK_WPT; e++) {
		const uint32_t L = len[e];
		if(!L) continue;
		uint32_t w[3];
		vi_encode(v[e], L, w);
		const uint32_t sh = 8 * fill;
		const uint32_t e0 = carry | (w[0] << sh), e1 = __funnelshift_l(w[0], w[1], sh),
			e2 = __funnelshift_l(w[1], w[2], sh), e3 = __funnelshift_l(w[2], 0u, sh);
		const uint32_t total = fill + L, nw = total >> 2;      // nw <= 3
		if(nw > 0) emit(e0);
		if(nw > 1) emit(e1);
		if(nw > 2) emit(e2);
		carry = nw == 0 ? e0 : nw == 1 ? e1 : nw == 2 ? e2 : e3;
		fill = total & 3u;
	}
	for(uint32_t b = partial ? head : 0; b < fill; b++) sb[wi * 4 + b] = (uint8_t) (carry >> (8 * b));
	__syncthreads();

	uint8_t *g = P.out + (uint64_t) p * P.stride + (goff - gshift);   // 16-byte aligned
	const uint32_t end = gshift + tile_bytes, chunks = (end + 15) >> 4;
	for(uint32_t c = threadIdx.x; c < chunks; c += WR_THREADS) {
		const uint32_t lo = c << 4;
		if(lo >= gshift && lo + 16 <= end) {
			__stcs(reinterpret_cast<uint4 *>(g + lo), *reinterpret_cast<const uint4 *>(sb + lo));
		} else {
			for(uint32_t b = max(lo, gshift); b < min(lo + 16, end); b++) g[b] = sb[b];
		}
	}
}

// ------------------------------------------------------------------ unpack

struct UnpackParams {
	const uint8_t *in;
	uint64_t stride;
	const uint64_t *sizes;
	uint64_t n;
	uint32_t tiles;
	void *vec;
	uint64_t ld;
	uint64_t *values;
	int32_t *status;
	uint32_t *tile_cnt;   // [count][tiles] terminator bytes of the payload inside a tile
	uint64_t *tile_off;   // [count][tiles] exclusive prefix = index of the first element starting in the tile
};

// header of message p: payload = bytes [*ps, *pe) of the message; returns false if malformed
__device__ __forceinline__ bool up_header(const uint8_t *m, uint64_t size, uint64_t *ps, uint64_t *pe) {
	*ps = *pe = 0;
	if(size == 0) return false;
	if(m[0] != 0x0A) return true;          // no vector field: trailer at 0
	uint64_t len = 0, o = 1;
	for(int k = 0; ; k++) {
		if(k == 10 || o >= size) return false;
		const uint8_t b = m[o++];
		if(7 * k < 64) len |= (uint64_t) (b & 0x7f) << (7 * k);
		if(!(b & 0x80)) break;
	}
	if(len > size - o) return false;
	*ps = o;
	*pe = o + len;
	return true;
}

// bit i of the result: byte i of the thread's 32 bytes has a clear MSB (ends a varint)
__device__ __forceinline__ uint32_t up_term_mask(const uint4 a, const uint4 b) {
	const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
	uint32_t m = 0;
#pragma unroll
	for(int i = 0; i < 8; i++) {
		// 0x80 in every terminator byte; bits 7, 15, 23, 31 gathered into a nibble by one multiplication
		// (the four partial products land on bits 21..24 and nothing else does)
		const uint32_t nib = ((((~w[i] & 0x80808080u) >> 7) * 0x00204081u) >> 21) & 0xfu;
		m |= nib << (4 * i);
	}
	return m;
}
// bits of a 32-byte window starting at message offset `base` that lie inside [ps, pe)
__device__ __forceinline__ uint32_t up_range_mask(uint64_t base, uint64_t ps, uint64_t pe) {
	if(pe <= base || ps >= base + 32) return 0;
	uint32_t m = 0xffffffffu;
	if(ps > base) m &= 0xffffffffu << (uint32_t) (ps - base);
	if(pe < base + 32) m &= 0xffffffffu >> (uint32_t) (base + 32 - pe);
	return m;
}

__global__ void __launch_bounds__(WR_THREADS) k_unpack_count(const UnpackParams P) {
	__shared__ uint32_t s_warp[WR_THREADS / 32 + 1];
	__shared__ uint64_t s_hdr[2];
	const uint32_t p = blockIdx.y, tile = blockIdx.x;
	const uint8_t *m = P.in + (uint64_t) p * P.stride;
	if(threadIdx.x == 0) {
		uint64_t ps, pe;
		const uint64_t size = P.sizes[p];
		if(size > P.stride || !up_header(m, size, &ps, &pe)) ps = pe = 0;
		s_hdr[0] = ps; s_hdr[1] = pe;
	}
	__syncthreads();
	const uint64_t ps = s_hdr[0], pe = s_hdr[1];
	const uint64_t base = (uint64_t) tile * UP_TILE + threadIdx.x * UP_BPT;
	uint32_t cnt = 0;
	const uint32_t rm = up_range_mask(base, ps, pe);
	if(rm) {
		const uint4 a = __ldg(reinterpret_cast<const uint4 *>(m + base));
		// the stride is a multiple of 16, not of 32: the second half of the last 32-byte group may lie past the buffer
		const uint4 b = base + 32 <= P.stride ? __ldg(reinterpret_cast<const uint4 *>(m + base + 16)) : make_uint4(0, 0, 0, 0);
		if(rm == 0xffffffffu) {
			cnt = __popc(~a.x & 0x80808080u) + __popc(~a.y & 0x80808080u) + __popc(~a.z & 0x80808080u) + __popc(~a.w & 0x80808080u)
				+ __popc(~b.x & 0x80808080u) + __popc(~b.y & 0x80808080u) + __popc(~b.z & 0x80808080u) + __popc(~b.w & 0x80808080u);
		} else cnt = __popc(up_term_mask(a, b) & rm);
	}
#pragma unroll
	for(int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
	if((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = cnt;
	__syncthreads();
	if(threadIdx.x == 0) {
		uint32_t t = 0;
		for(int i = 0; i < WR_THREADS / 32; i++) t += s_warp[i];
		P.tile_cnt[(uint64_t) p * P.tiles + tile] = t;
	}
}

// one CTA per message: tile offsets, element count, trailer [0x10][value], status
__global__ void __launch_bounds__(WR_THREADS) k_unpack_scan(const UnpackParams P, int ring_bits) {
	__shared__ uint64_t s_part[WR_THREADS];
	__shared__ uint64_t s_total;
	const uint32_t p = blockIdx.x, t = threadIdx.x;
	const uint32_t per = (P.tiles + WR_THREADS - 1) / WR_THREADS;
	const uint32_t lo = min(t * per, P.tiles), hi = min(lo + per, P.tiles);
	const uint32_t *tc = P.tile_cnt + (uint64_t) p * P.tiles;
	uint64_t *to = P.tile_off + (uint64_t) p * P.tiles;
	uint64_t sum = 0;
	for(uint32_t i = lo; i < hi; i++) sum += tc[i];
	s_part[t] = sum;
	__syncthreads();
	if(t == 0) {
		uint64_t run = 0;
		for(int i = 0; i < WR_THREADS; i++) { const uint64_t c = s_part[i]; s_part[i] = run; run += c; }
		s_total = run;
	}
	__syncthreads();
	uint64_t run = s_part[t];
	for(uint32_t i = lo; i < hi; i++) { to[i] = run; run += tc[i]; }
	if(t == 0) {
		const uint8_t *m = P.in + (uint64_t) p * P.stride;
		const uint64_t size = P.sizes[p];
		int32_t st = 0;
		uint64_t ps = 0, pe = 0, val = 0;
		if(size > P.stride || !up_header(m, size, &ps, &pe)) st |= P1_MSG_BAD_HEADER;
		else {
			if(s_total != P.n) st |= P1_MSG_BAD_COUNT;
			if(pe > ps && (m[pe - 1] & 0x80)) st |= P1_MSG_BAD_VARINT;     // payload ends inside an element
			uint64_t o = pe;
			if(o >= size || m[o] != 0x10) st |= P1_MSG_BAD_HEADER;
			else {
				o++;
				bool done = false;
				for(int k = 0; k < 10 && o < size; k++) {
					const uint8_t b = m[o++];
					if(7 * k < 64) val |= (uint64_t) (b & 0x7f) << (7 * k);
					if(!(b & 0x80)) { done = true; break; }
				}
				if(!done || o != size) st |= P1_MSG_BAD_HEADER;
			}
		}
		if(ring_bits == 32) val &= 0xffffffffull;
		P.values[p] = st ? 0 : val;
		P.status[p] = st;
	}
}

// every byte that follows a terminator (or opens the payload) starts an element; its index
// is the number of terminators before it
template <typename W>
__global__ void __launch_bounds__(WR_THREADS) k_unpack_write(const UnpackParams P) {
	__shared__ __align__(16) uint8_t sb[16 + UP_TILE + 16];   // 16 bytes of halo on both sides
	__shared__ uint32_t s_warp[WR_THREADS / 32 + 1];
	__shared__ uint64_t s_hdr[2];
	const uint32_t p = blockIdx.y, tile = blockIdx.x;
	const uint8_t *m = P.in + (uint64_t) p * P.stride;
	if(threadIdx.x == 0) {
		uint64_t ps, pe;
		const uint64_t size = P.sizes[p];
		if(size > P.stride || !up_header(m, size, &ps, &pe)) ps = pe = 0;
		s_hdr[0] = ps; s_hdr[1] = pe;
	}
	__syncthreads();
	const uint64_t ps = s_hdr[0], pe = s_hdr[1];
	const uint64_t t0 = (uint64_t) tile * UP_TILE;
	if(pe <= t0 || ps >= t0 + UP_TILE) return;      // no payload byte in this tile (uniform)
	// stage [t0 - 16, t0 + UP_TILE + 16), zero outside the message's stride
	for(uint32_t c = threadIdx.x; c < (UP_TILE + 32) / 16; c += WR_THREADS) {
		const int64_t off = (int64_t) t0 - 16 + (int64_t) c * 16;
		uint4 q = make_uint4(0, 0, 0, 0);
		if(off >= 0 && (uint64_t) off + 16 <= P.stride) q = __ldcs(reinterpret_cast<const uint4 *>(m + off));
		*reinterpret_cast<uint4 *>(sb + c * 16) = q;
	}
	__syncthreads();
	const uint32_t lb = 16 + threadIdx.x * UP_BPT;      // the thread's first byte in sb
	const uint64_t base = t0 + threadIdx.x * UP_BPT;    // ... and in the message
	const uint4 a = *reinterpret_cast<const uint4 *>(sb + lb), b = *reinterpret_cast<const uint4 *>(sb + lb + 16);
	const uint32_t rm = up_range_mask(base, ps, pe);
	const uint32_t term = up_term_mask(a, b);
	const uint32_t prev_term = !(sb[lb - 1] & 0x80);   // the byte before pay_start is the last byte of the length varint
	const uint32_t starts = ((term << 1) | prev_term) & rm;
	uint32_t tile_total;
	const uint32_t before = p1_block_excl_scan256(__popc(term & rm), s_warp, &tile_total);
	uint64_t index = P.tile_off[(uint64_t) p * P.tiles + tile] + before;
	W *out = (W *) P.vec + (uint64_t) p * P.ld;
	uint32_t bad = 0;
	for(uint32_t s = starts; s; s &= s - 1) {
		const uint32_t i = __ffs(s) - 1;
		const uint64_t k = index + __popc(term & rm & ((1u << i) - 1u));
		// 12 bytes from the aligned word below the start, shifted down to the start
		const uint32_t o = lb + i, sh = 8 * (o & 3u);
		const uint32_t *wp = reinterpret_cast<const uint32_t *>(sb + (o & ~3u));
		const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
		const uint32_t d0 = __funnelshift_r(w0, w1, sh), d1 = __funnelshift_r(w1, w2, sh), d2 = __funnelshift_r(w2, w3, sh);
		// all ten candidate bytes decoded, then cut at the first terminator
		uint64_t v = (uint64_t) vi_gather28(d0 & 0x7f7f7f7fu) | ((uint64_t) vi_gather28(d1 & 0x7f7f7f7fu) << 28)
			| ((uint64_t) (d2 & 0x7fu) << 56) | ((uint64_t) ((d2 >> 8) & 1u) << 63);
		const uint32_t t0m = ~d0 & 0x80808080u, t1m = ~d1 & 0x80808080u;
		uint32_t len;
		if(t0m) len = __ffs(t0m) >> 3;
		else if(t1m) len = 4 + (__ffs(t1m) >> 3);
		else if(!(d2 & 0x80u)) len = 9;
		else if(!(d2 & 0x8000u)) len = 10;
		else { len = 10; bad = 1; }
		if(len < 10) v &= (1ull << (7 * len)) - 1;
		if(k < P.n) out[k] = (W) v;
	}
	if(bad) atomicOr(P.status + p, P1_MSG_BAD_VARINT);
}

// ------------------------------------------------------------------ host side

static int wire_ws(p1_ctx *ctx, size_t count, size_t tiles, uint32_t **sum, uint64_t **off) {
	const size_t ent = count * (tiles ? tiles : 1);
	const size_t need = p1_round_up(ent * 8, 256) + p1_round_up(ent * 4, 256);
	if(ctx->wire_ws_bytes < need) {
		if(ctx->wire_ws) {
			P1_CUDA(cudaDeviceSynchronize());
			p1_pool_dev_give(ctx->device, ctx->wire_ws, ctx->wire_ws_bytes);
			ctx->wire_ws_bytes = 0;
		}
		ctx->wire_ws = p1_pool_dev_take(ctx->device, need);
		P1_CHECK(ctx->wire_ws, "p1_msg: cannot allocate %zu bytes of workspace", need);
		ctx->wire_ws_bytes = need;
	}
	*off = (uint64_t *) ctx->wire_ws;
	*sum = (uint32_t *) ((char *) ctx->wire_ws + p1_round_up(ent * 8, 256));
	return 0;
}

extern "C" size_t p1_msg_max_bytes(const p1_ctx *ctx, size_t n_words) {
	if(!ctx) return 0;
	const size_t per = ctx->bits_p1 == 64 ? 10 : 5;
	// tag + length varint + elements + tag + value, rounded up so that it can serve as a stride
	return p1_round_up(1 + 10 + n_words * per + 1 + per, 16);
}

extern "C" int p1_msg_pack(p1_ctx *ctx, const void *vec_dev, size_t ld_words, size_t n_words, size_t count,
	const uint64_t *values, uint8_t *out_dev, size_t stride_bytes, uint64_t *sizes_dev) {
	P1_ENTER(ctx);
	if(!count) return 0;
	const size_t w = ctx->bits_p1 / 8;
	P1_CHECK(values && out_dev && sizes_dev && (vec_dev || !n_words), "p1_msg_pack: arguments may not be null");
	P1_CHECK(stride_bytes % 16 == 0 && stride_bytes >= p1_msg_max_bytes(ctx, n_words),
		"p1_msg_pack: stride of %zu bytes (need a multiple of 16, at least %zu)", stride_bytes, p1_msg_max_bytes(ctx, n_words));
	P1_CHECK(((uintptr_t) out_dev & 15) == 0 && ((uintptr_t) vec_dev & 15) == 0 && (ld_words * w) % 16 == 0 && (count == 1 || ld_words >= n_words),
		"p1_msg_pack: buffers must be 16-byte aligned and vectors at least n_words apart");
	P1_CHECK(count <= 65535, "p1_msg_pack: at most 65535 messages per call");
	const size_t tiles = (n_words + PK_TILE - 1) / PK_TILE;
	P1_CHECK(tiles < (1u << 31), "p1_msg_pack: vector too long");
	PackParams P;
	P.vec = vec_dev; P.ld = ld_words; P.n = n_words; P.tiles = (uint32_t) tiles;
	P.out = out_dev; P.stride = stride_bytes; P.sizes = sizes_dev;
	const void *vd = NULL;
	if(p1_small_to_dev(ctx, values, count * sizeof(uint64_t), 0, &vd)) return 1;
	P.values = (const uint64_t *) vd;
	if(wire_ws(ctx, count, tiles, &P.tile_sum, &P.tile_off)) return 1;
	const dim3 grid((unsigned) tiles, (unsigned) count);
	p1_prof_begin(ctx, P1_K_MSG_PACK);
	if(tiles) {
		if(w == 8) k_pack_len<uint64_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		else k_pack_len<uint32_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		ctx->launches++;
	}
	k_pack_scan<<<(unsigned) count, WR_THREADS, 0, ctx->stream>>>(P, ctx->bits_p1);
	if(tiles) {
		ctx->launches++;
		if(w == 8) k_pack_write<uint64_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		else k_pack_write<uint32_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
	}
	P1_LAUNCHED(ctx);
	return 0;
}

// ------------------------------------------------------------------ p1_msg_pack_static: the same pack, no allocation
//
// p1_msg_pack needs 12 bytes of device scratch per 2048-word tile (k_pack_len's byte counts, k_pack_scan's
// tile offsets) and takes them from the block pool, i.e. from cudaMalloc when the pool has no block of that
// size.  Some hosts of this library cannot afford that call at the moment they send a message:
//   - a provider process that gave all of the GPU's memory to its own arena at start-up (X, y and the message
//     buffers of a 180 GB shard are sized once; nothing is left for the driver to hand out later);
//   - the per-peer threads of --use_ot mode, which share one GPU: cudaMalloc and cudaFree wait for the whole
//     device, so one thread's allocation stalls the kernels of every other peer's session;
//   - a party run under an operator-set memory cap, where an allocation in the middle of the protocol may be
//     refused after the peers have already been sent half of their messages.
// For them the scratch lives in the library image itself: the array below is part of libp1gpu.so's device
// data, so every GPU on which the library's kernels are loaded holds these 768 KB until the process exits,
// whether this entry point is used or not.  The three kernels are the ones p1_msg_pack launches; only where
// their tile tables live differs, so the bytes produced are the same.
//
// In host code `g_pack_static` is only a handle for the runtime - not an address a kernel could use.  The
// real address differs per GPU and is looked up with cudaGetSymbolAddress the first time the entry point runs
// on a device (with that device current), then remembered.
static __device__ __align__(256) unsigned char g_pack_static[(size_t) P1_MSG_PACK_STATIC_MAX_TILES * 12];

#define PK_STATIC_MAX_DEVICES 64
static std::mutex g_pack_static_mu;
static void *g_pack_static_addr[PK_STATIC_MAX_DEVICES];   // per CUDA device ordinal, NULL = not looked up yet

// device address of g_pack_static on the current device (the caller has made `device` current)
static int pack_static_addr(int device, void **addr) {
	P1_CHECK(device >= 0 && device < PK_STATIC_MAX_DEVICES, "p1_msg_pack_static: device ordinal %d not supported (max %d)",
		device, PK_STATIC_MAX_DEVICES - 1);
	std::lock_guard<std::mutex> l(g_pack_static_mu);
	if(!g_pack_static_addr[device]) {
		void *p = NULL;
		P1_CUDA(cudaGetSymbolAddress(&p, g_pack_static));
		P1_CHECK(p, "p1_msg_pack_static: no device address for the scratch array on device %d", device);
		g_pack_static_addr[device] = p;
	}
	*addr = g_pack_static_addr[device];
	return 0;
}

extern "C" int p1_msg_pack_static(p1_ctx *ctx, const void *vec_dev, size_t ld_words, size_t n_words, size_t count,
	const uint64_t *values_dev, uint8_t *out_dev, size_t stride_bytes, uint64_t *sizes_dev) {
	P1_ENTER(ctx);
	if(!count) return 0;
	const size_t w = ctx->bits_p1 / 8;
	P1_CHECK(values_dev && out_dev && sizes_dev && (vec_dev || !n_words), "p1_msg_pack_static: arguments may not be null");
	P1_CHECK(count <= 65535, "p1_msg_pack_static: at most 65535 messages per call");
	// size limit first: a call that is too large is refused before any pointer is looked at
	const size_t tiles = (n_words + PK_TILE - 1) / PK_TILE;
	const size_t ent = count * (tiles ? tiles : 1);
	P1_CHECK(tiles <= P1_MSG_PACK_STATIC_MAX_TILES && ent <= P1_MSG_PACK_STATIC_MAX_TILES,
		"p1_msg_pack_static: %zu messages of %zu words need %zu tile entries, the built-in scratch holds %d "
		"(count * ceil(n_words / %d) <= %d); split the batch or use p1_msg_pack",
		count, n_words, ent, P1_MSG_PACK_STATIC_MAX_TILES, PK_TILE, P1_MSG_PACK_STATIC_MAX_TILES);
	P1_CHECK(stride_bytes % 16 == 0 && stride_bytes >= p1_msg_max_bytes(ctx, n_words),
		"p1_msg_pack_static: stride of %zu bytes (need a multiple of 16, at least %zu)", stride_bytes, p1_msg_max_bytes(ctx, n_words));
	P1_CHECK(((uintptr_t) out_dev & 15) == 0 && ((uintptr_t) vec_dev & 15) == 0 && (ld_words * w) % 16 == 0 && (count == 1 || ld_words >= n_words),
		"p1_msg_pack_static: buffers must be 16-byte aligned and vectors at least n_words apart");
	// p1_msg_pack stages a host `values` array through the context's copy slot; here nothing is staged
	P1_CHECK(p1_is_device_ptr(values_dev), "p1_msg_pack_static: values must already be on the device");
	void *ws = NULL;
	if(pack_static_addr(ctx->device, &ws)) return 1;
	PackParams P;
	P.vec = vec_dev; P.ld = ld_words; P.n = n_words; P.tiles = (uint32_t) tiles;
	P.out = out_dev; P.stride = stride_bytes; P.sizes = sizes_dev; P.values = values_dev;
	// same split as wire_ws: offsets first, sums behind them.  Nothing has to start at zero: k_pack_len writes
	// every tile_sum entry of this call before k_pack_scan reads it, and k_pack_scan writes every tile_off entry
	// before k_pack_write reads it, so what an earlier call left in the array is never seen.
	P.tile_off = (uint64_t *) ws;
	P.tile_sum = (uint32_t *) ((char *) ws + p1_round_up(ent * 8, 256));
	const dim3 grid((unsigned) tiles, (unsigned) count);
	if(tiles) {
		if(w == 8) k_pack_len<uint64_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		else k_pack_len<uint32_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		ctx->launches++;
	}
	k_pack_scan<<<(unsigned) count, WR_THREADS, 0, ctx->stream>>>(P, ctx->bits_p1);
	ctx->launches++;
	if(tiles) {
		if(w == 8) k_pack_write<uint64_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		else k_pack_write<uint32_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
		ctx->launches++;
	}
	const cudaError_t e = cudaGetLastError();
	P1_CHECK(e == cudaSuccess, "p1_msg_pack_static: kernel launch: %s", cudaGetErrorString(e));
	return 0;
}

extern "C" int p1_msg_unpack(p1_ctx *ctx, const uint8_t *in_dev, size_t stride_bytes, const uint64_t *sizes, size_t count,
	size_t n_words, void *vec_dev, size_t ld_words, uint64_t *values_dev, int32_t *status_dev) {
	P1_ENTER(ctx);
	if(!count) return 0;
	const size_t w = ctx->bits_p1 / 8;
	P1_CHECK(in_dev && sizes && values_dev && status_dev && (vec_dev || !n_words), "p1_msg_unpack: arguments may not be null");
	P1_CHECK(stride_bytes % 16 == 0 && stride_bytes >= 16 && ((uintptr_t) in_dev & 15) == 0,
		"p1_msg_unpack: the input must be 16-byte aligned with a stride that is a multiple of 16");
	P1_CHECK(count == 1 || ld_words >= n_words, "p1_msg_unpack: ld_words %zu < n_words %zu", ld_words, n_words);
	P1_CHECK(count <= 65535, "p1_msg_unpack: at most 65535 messages per call");
	const size_t tiles = (stride_bytes + UP_TILE - 1) / UP_TILE;
	P1_CHECK(tiles < (1u << 31), "p1_msg_unpack: message too long");
	UnpackParams P;
	P.in = in_dev; P.stride = stride_bytes; P.n = n_words; P.tiles = (uint32_t) tiles;
	P.vec = vec_dev; P.ld = ld_words; P.values = values_dev; P.status = status_dev;
	const void *sd = NULL;
	if(p1_small_to_dev(ctx, sizes, count * sizeof(uint64_t), 0, &sd)) return 1;
	P.sizes = (const uint64_t *) sd;
	if(wire_ws(ctx, count, tiles, &P.tile_cnt, &P.tile_off)) return 1;
	const dim3 grid((unsigned) tiles, (unsigned) count);
	p1_prof_begin(ctx, P1_K_MSG_UNPACK);
	k_unpack_count<<<grid, WR_THREADS, 0, ctx->stream>>>(P);
	ctx->launches++;
	k_unpack_scan<<<(unsigned) count, WR_THREADS, 0, ctx->stream>>>(P, ctx->bits_p1);
	ctx->launches++;
	if(w == 8) k_unpack_write<uint64_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
	else k_unpack_write<uint32_t><<<grid, WR_THREADS, 0, ctx->stream>>>(P);
	P1_LAUNCHED(ctx);
	return 0;
}

// ------------------------------------------------------------------ f3: choice bits of the share hand-off (src/input.c:41)

template <typename W>
__global__ void __launch_bounds__(256) k_share_bits(const W *__restrict__ in, size_t n, uint8_t *__restrict__ sel) {
	constexpr int B = sizeof(W) * 8, Q = B / 16;      // 16-byte stores per word
	const size_t total = n * Q;
	for(size_t t = blockIdx.x * (size_t) blockDim.x + threadIdx.x; t < total; t += (size_t) gridDim.x * blockDim.x) {
		const size_t i = t / Q;
		const uint32_t h = (uint32_t) ((in[i] >> (16 * (t % Q))) & 0xffffu);
		uint32_t o[4];
#pragma unroll
		for(int c = 0; c < 4; c++) {
			const uint32_t nib = (h >> (4 * c)) & 0xfu;
			o[c] = (nib & 1u) | ((nib & 2u) << 7) | ((nib & 4u) << 14) | ((nib & 8u) << 21);
		}
		reinterpret_cast<uint4 *>(sel)[t] = make_uint4(o[0], o[1], o[2], o[3]);
	}
}

extern "C" int p1_share_bits(p1_ctx *ctx, const void *shares, size_t n, uint8_t *sel) {
	P1_ENTER(ctx);
	if(!n) return 0;
	P1_CHECK(shares && sel, "p1_share_bits: arguments may not be null");
	const size_t w = ctx->bits_p2 / 8, in_bytes = n * w, out_bytes = n * ctx->bits_p2;
	const bool in_dev = p1_is_device_ptr(shares), out_dev = p1_is_device_ptr(sel);
	p1_tmp scratch;
	void *tmp = NULL;
	const size_t tmp_bytes = (in_dev ? 0 : p1_round_up(in_bytes, 256)) + (out_dev ? 0 : out_bytes);
	if(tmp_bytes) {
		P1_CUDA(scratch.alloc(tmp_bytes, ctx->stream));
		tmp = scratch.p;
	}
	const void *src = shares;
	uint8_t *dst = sel;
	if(!in_dev) {
		P1_CUDA(cudaMemcpyAsync(tmp, shares, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
		src = tmp;
	}
	if(!out_dev) dst = (uint8_t *) tmp + (in_dev ? 0 : p1_round_up(in_bytes, 256));
	P1_CHECK(((uintptr_t) dst & 15) == 0 && ((uintptr_t) src & (w - 1)) == 0, "p1_share_bits: sel must be 16-byte aligned");
	const size_t threads = n * (ctx->bits_p2 / 16);
	const unsigned grid = (unsigned) ((threads + 255) / 256 < (size_t) ctx->sm_count * 8 ? (threads + 255) / 256 : (size_t) ctx->sm_count * 8);
	p1_prof_begin(ctx, P1_K_SHARE_BITS);
	if(w == 8) k_share_bits<uint64_t><<<grid, 256, 0, ctx->stream>>>((const uint64_t *) src, n, dst);
	else k_share_bits<uint32_t><<<grid, 256, 0, ctx->stream>>>((const uint32_t *) src, n, dst);
	P1_LAUNCHED(ctx);
	if(!out_dev) {
		P1_CUDA(cudaMemcpyAsync(sel, dst, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
		P1_CUDA(cudaStreamSynchronize(ctx->stream));
	}
	return 0;
}
