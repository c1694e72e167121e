// FASTA text -> encoded sequence on the device (for the index path).
//
// The reference turns FASTA into symbol codes one character at a time (gt_sequence_buffer_fasta_advance,
// src/core/sequence_buffer_fasta.c:40-170, called from the fill loops of gt_encseq_encoder_encode,
// src/core/encseq.c:8523): outside a description every non-space character is a symbol, except '>' which starts a
// description (and, but for the first one, emits a separator); a description ends at the next newline.  The state
// "inside a description" after a byte is therefore decided by the last '>' or newline at or before it - which makes
// the scan parallel:
//   k_fa_summary  one thread per 256-byte chunk: does the chunk contain '>' or '\n', and which comes last
//   (exclusive scan of the "contains" flags, k_fa_compact)  -> the state at the start of every chunk
//   k_fa_count    symbols and description starts per chunk (and the first illegal character, if any)
//   (exclusive scans of both counts)
//   k_fa_emit     writes code(symbol) at  #symbols before + #descriptions before - 1, and a separator for every
//                 description but the first
// The result is the byte form gtb200_open takes (codes 0..sigma-1, 254 wildcard, 255 separator); packing to 2 bits,
// the special bitmap and the range table are the kernels gtb200_open already runs.  Symbol maps: src/core/alphabet.c:84-91.
// Included at the end of gtsfx_b200.cu.
#pragma once

#define FA_CHUNK 256

__device__ __forceinline__ uint8_t fa_code_dna(uint8_t ch) {
  switch (ch | 0x20) {               // ASCII lower case
    case 'a': return 0;
    case 'c': return 1;
    case 'g': return 2;
    case 't': case 'u': return 3;
    case 'n': case 's': case 'y': case 'w': case 'r': case 'k': case 'v': case 'b': case 'd': case 'h': case 'm': return 254;
    default: return 253;             // not a character of the alphabet
  }
}

__device__ __forceinline__ uint8_t fa_code_protein(uint8_t ch) {
  // LVIFKREDAGSTNQYWPHMC -> 0..19, XUBZJO*- -> wildcard (upper case only, like the reference's map)
  switch (ch) {
    case 'L': return 0;  case 'V': return 1;  case 'I': return 2;  case 'F': return 3;  case 'K': return 4;
    case 'R': return 5;  case 'E': return 6;  case 'D': return 7;  case 'A': return 8;  case 'G': return 9;
    case 'S': return 10; case 'T': return 11; case 'N': return 12; case 'Q': return 13; case 'Y': return 14;
    case 'W': return 15; case 'P': return 16; case 'H': return 17; case 'M': return 18; case 'C': return 19;
    case 'X': case 'U': case 'B': case 'Z': case 'J': case 'O': case '*': case '-': return 254;
    default: return 253;
  }
}

// isspace() of the C locale
__device__ __forceinline__ bool fa_isspace(uint8_t ch) { return ch == ' ' || (ch >= 9 && ch <= 13); }

// f is the device copy of the text, padded with blanks to a multiple of FA_CHUNK (a blank is ignored in either
// state), so a chunk is 16 aligned 16-byte loads without bounds checks
__global__ void __launch_bounds__(256) k_fa_summary(const uint8_t *__restrict__ f, uint64_t nchunks,
                                                    uint32_t *__restrict__ det, uint8_t *__restrict__ val) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nchunks) return;
  const uint4 *f4 = reinterpret_cast<const uint4 *>(f + t * FA_CHUNK);
  uint32_t d = 0;
  uint8_t v = 0;
#pragma unroll 4
  for (int q = 0; q < FA_CHUNK / 16; ++q) {
    const uint4 x = __ldg(f4 + q);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const uint8_t ch = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
      if (ch == '>') { d = 1; v = 1; }
      else if (ch == '\n') { d = 1; v = 0; }
    }
  }
  det[t] = d;
  val[t] = v;
}

__global__ void __launch_bounds__(256) k_fa_compact(const uint32_t *__restrict__ det, const uint32_t *__restrict__ detx,
                                                    const uint8_t *__restrict__ val, uint64_t nchunks,
                                                    uint8_t *__restrict__ vals) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nchunks && det[t]) vals[detx[t]] = val[t];
}

// PASS 0: symbols and description starts per chunk, first illegal character.  PASS 1: emit (nsym/nhdr hold the
// exclusive prefix sums by then).
template <int PASS, bool PROTEIN>
__global__ void __launch_bounds__(256) k_fa_walk(const uint8_t *__restrict__ f, uint64_t nchunks,
                                                 const uint32_t *__restrict__ detx, const uint8_t *__restrict__ vals,
                                                 uint32_t *__restrict__ nsym, uint32_t *__restrict__ nhdr,
                                                 unsigned long long *__restrict__ firstbad,
                                                 uint8_t *__restrict__ enc) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nchunks) return;
  const uint4 *f4 = reinterpret_cast<const uint4 *>(f + t * FA_CHUNK);
  bool indesc = detx[t] ? vals[detx[t] - 1] != 0 : false;     // state after the last '>' / newline before the chunk
  uint32_t s = 0, h = 0;
  uint64_t spos = 0, hcount = 0;
  if (PASS == 1) { spos = nsym[t]; hcount = nhdr[t]; }
#pragma unroll 2
  for (int q = 0; q < FA_CHUNK / 16; ++q) {
    const uint4 x = __ldg(f4 + q);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const uint8_t ch = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
      if (indesc) {
        if (ch == '\n') indesc = false;
      } else if (ch == '>') {
        indesc = true;
        if (PASS == 0) ++h;
        else {
          if (hcount > 0) enc[spos + hcount - 1] = 255;        // separator before every sequence but the first
          ++hcount;
        }
      } else if (!fa_isspace(ch)) {
        const uint8_t code = PROTEIN ? fa_code_protein(ch) : fa_code_dna(ch);
        if (PASS == 0) {
          if (code == 253) atomicMin(firstbad, (unsigned long long)(t * FA_CHUNK + q * 16 + j));
          ++s;
        } else {
          enc[spos + (hcount ? hcount - 1 : 0)] = code;
          ++spos;
        }
      }
    }
  }
  if (PASS == 0) { nsym[t] = s; nhdr[t] = h; }
}

// ---- host side ----
extern "C" unsigned gtb200_recommendedprefixlength(unsigned numofchars, uint64_t totallength);

extern "C" int gtb200_open_fasta(gtb200_ctx **out, int device, const uint8_t *fasta, uint64_t nbytes, int protein,
                                 unsigned prefixlength, gtb200_fastainfo *info) {
  if (!out || (!fasta && nbytes)) return set_error("gtb200_open_fasta: null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return set_error("gtb200_open_fasta: no CUDA device available; this library has no CPU fallback");
  if (device < 0 || device >= ndev) return set_error("gtb200_open_fasta: device %d out of range (0..%d)", device, ndev - 1);
  if (nbytes >= (1ull << 40)) return set_error("gtb200_open_fasta: %llu bytes of text are too many", (unsigned long long)nbytes);
  CK(cudaSetDevice(device));
  const uint64_t nchunks = (nbytes + FA_CHUNK - 1) / FA_CHUNK + 1;      // at least one, the last all blanks
  // a scratch context: device allocations come from (and go back to) the library's block cache
  gtb200_ctx *tmp = ctx_create(device, protein ? 20 : 4, 0, 1);
  if (!tmp) return -1;
  auto fail = [&](int rc) { free_ctx(tmp); return rc; };
  uint8_t *d_f = nullptr, *d_val = nullptr, *d_vals = nullptr, *d_enc = nullptr;
  uint32_t *d_det = nullptr, *d_detx = nullptr, *d_nsym = nullptr, *d_nhdr = nullptr;
  unsigned long long *d_tot = nullptr;     // [0] symbols, [1] descriptions, [2] first illegal character, [3] determined chunks
  if (dmalloc(tmp, &d_f, nchunks * FA_CHUNK) != cudaSuccess || dmalloc(tmp, &d_val, nchunks) != cudaSuccess ||
      dmalloc(tmp, &d_vals, nchunks + 1) != cudaSuccess || dmalloc(tmp, &d_det, (nchunks + 1) * 4) != cudaSuccess ||
      dmalloc(tmp, &d_detx, (nchunks + 1) * 4) != cudaSuccess || dmalloc(tmp, &d_nsym, (nchunks + 1) * 4) != cudaSuccess ||
      dmalloc(tmp, &d_nhdr, (nchunks + 1) * 4) != cudaSuccess || dmalloc(tmp, &d_tot, 64) != cudaSuccess) {
    set_error("gtb200_open_fasta: out of device memory");
    return fail(-1);
  }
  cudaStream_t st = tmp->stream;
  cudaError_t e = cudaMemsetAsync(d_f + (nbytes / FA_CHUNK) * FA_CHUNK, ' ', nchunks * FA_CHUNK - (nbytes / FA_CHUNK) * FA_CHUNK, st);
  if (e == cudaSuccess && nbytes) e = cudaMemcpyAsync(d_f, fasta, nbytes, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_tot, 0xff, 32, st);
  if (e != cudaSuccess) { set_error("gtb200_open_fasta: upload failed: %s", cudaGetErrorString(e)); return fail(-1); }
  const unsigned grid = grid_for(nchunks, 256);
  {
    ProfScope ps(tmp, KC_PREP, 1);
    k_fa_summary<<<grid, 256, 0, st>>>(d_f, nchunks, d_det, d_val);
  }
  if (device_scan(tmp, d_det, d_detx, nchunks, 0, d_tot + 3, KC_PREP)) return fail(-1);
  {
    ProfScope ps(tmp, KC_PREP, 2);
    k_fa_compact<<<grid, 256, 0, st>>>(d_det, d_detx, d_val, nchunks, d_vals);
    if (protein) k_fa_walk<0, true><<<grid, 256, 0, st>>>(d_f, nchunks, d_detx, d_vals, d_nsym, d_nhdr, d_tot + 2, nullptr);
    else k_fa_walk<0, false><<<grid, 256, 0, st>>>(d_f, nchunks, d_detx, d_vals, d_nsym, d_nhdr, d_tot + 2, nullptr);
  }
  if (device_scan(tmp, d_nsym, d_nsym, nchunks, 0, d_tot + 0, KC_PREP)) return fail(-1);
  if (device_scan(tmp, d_nhdr, d_nhdr, nchunks, 0, d_tot + 1, KC_PREP)) return fail(-1);
  unsigned long long h_tot[4];
  e = cudaMemcpyAsync(h_tot, d_tot, 32, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { set_error("gtb200_open_fasta: %s", cudaGetErrorString(e)); return fail(-1); }
  if (h_tot[2] != ~0ull) {
    set_error("gtb200_open_fasta: illegal character at byte %llu of the input (not a symbol of the %s alphabet)", h_tot[2],
              protein ? "protein" : "DNA");
    return fail(-1);
  }
  if (h_tot[1] == 0) { set_error("gtb200_open_fasta: no sequences in the input (no '>' found)"); return fail(-1); }
  const uint64_t totallength = h_tot[0] + h_tot[1] - 1;
  if (totallength >= 0xfffffffeull) {
    set_error("gtb200_open_fasta: totallength %llu does not fit the 32-bit bucket table", (unsigned long long)totallength);
    return fail(-1);
  }
  if (dmalloc(tmp, &d_enc, totallength + 1) != cudaSuccess) { set_error("gtb200_open_fasta: out of device memory"); return fail(-1); }
  {
    ProfScope ps(tmp, KC_PREP, 1);
    if (protein) k_fa_walk<1, true><<<grid, 256, 0, st>>>(d_f, nchunks, d_detx, d_vals, d_nsym, d_nhdr, nullptr, d_enc);
    else k_fa_walk<1, false><<<grid, 256, 0, st>>>(d_f, nchunks, d_detx, d_vals, d_nsym, d_nhdr, nullptr, d_enc);
  }
  e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { set_error("gtb200_open_fasta: %s", cudaGetErrorString(e)); return fail(-1); }
  gtb200_seqdesc seq;
  memset(&seq, 0, sizeof seq);
  seq.numofchars = protein ? 20 : 4;
  seq.totallength = totallength;
  seq.encoded = d_enc;
  seq.memory = GTB200_MEM_DEVICE;
  if (prefixlength == 0) prefixlength = gtb200_recommendedprefixlength(seq.numofchars, totallength);
  const int rc = gtb200_open(out, device, &seq, prefixlength);
  if (rc == 0 && info) {
    info->totallength = totallength;
    info->numofsequences = h_tot[1];
    info->prefixlength = prefixlength;
    info->specialcharacters = (*out)->specialchars;
  }
  return fail(rc);       // releases the scratch context (its blocks return to the cache) in either case
}
