// Group keys wider than 64 packed bits, or more than 4 key columns (included by groupby.cu in its anonymous namespace).
//
// The reference's Grouper takes any number of fixed-width key columns: KeyEncoder::Encode writes, per key column,
// [1 null byte][the value's bytes, zero when null] into a row-major byte string and the string is the hash-map key
// (hash_aggregate.cc:48-196, 314-408).  The aggregation kernels of this library work on a key packed into one 64-bit
// register, so a wide key is first COMPRESSED: a second open-addressing table maps the encoded key (the same byte
// string as upstream's, held in up to kWideWords 64-bit words) to a dense 32-bit id, the rows' ids form an ordinary
// uint32 key column, and everything downstream (all consume paths, the grouper, finalize) runs on that column.  At
// finalize the ids are decoded back into the key columns from the key store.
//   slots   uint64 per slot: EMPTY, or {kind, 31-bit hash tag, 32-bit payload}: kind 0 = the key of row `payload` of
//           the batch being inserted (claimed with ONE atomicCAS, no lock: the key bytes are read from the batch's own
//           columns), kind 1 = the key stored as id `payload` in the key store (every earlier batch)
//   store   id -> the encoded key words
// Three passes per batch over the key columns: insert (claim / find a slot per row), assign (new slots get their ids,
// keys copied to the store), ids (row -> id).
//
// VARIABLE-LENGTH key columns (binary / utf8 with int32 offsets; VarLengthKeyEncoder, hash_aggregate.cc:236-311) ride on
// the same table.  Their part of the encoded key is [null byte][4-byte length]; the bytes themselves are hashed and
// compared where they lie — in the batch's value buffer for a row of the batch, in a byte heap for a stored key —
// and the heap position of every stored key's bytes is kept beside the store (`store_var`).  New keys reach the heap
// through one scan of their lengths.

constexpr int kWideMaxKeys = 16;
constexpr int kWideWords = 18;   // 16 keys x (1 + 8) bytes = 144 bytes
constexpr uint64_t kWideEmpty = ~uint64_t(0);
constexpr int kWideProbe = 1 << 20;

struct WideKeyCol {
  const void* data;   // shifted by the element offset (non-bool)
  BitmapRef bits;     // bool key data
  BitmapRef vv;
  int32_t nbits;      // 0: variable-length (offsets / bytes below)
  int32_t byte_off;   // where this column's [null byte][value bytes] start in the encoded key
  const int32_t* offsets;   // variable-length: shifted by the element offset
  const uint8_t* bytes;
  int32_t var_index;        // position among the variable-length key columns, -1 for a fixed-width one
  int32_t pad;
};

struct WideArgs {
  WideKeyCol cols[kWideMaxKeys];
  int32_t n_keys;
  int32_t words;        // 64-bit words of the encoded key
  int64_t n;
  uint64_t* slots;
  uint64_t cap;         // power of two
  uint64_t* store;      // [ids][words]
  uint32_t* counters;   // [0] ids in the store, [1] slots claimed by this batch, [2] error
  uint32_t* new_slots;  // slot indices claimed by this batch
  uint32_t* row_slot;   // per row: its slot
  uint32_t* ids;        // per row: its id
  const uint32_t* select;   // nullable: rows whose bit is clear get no slot (their id is never looked at)
  // variable-length key columns
  int32_t n_var;
  int32_t var_col[kWideMaxKeys];   // key column of var_index v
  uint64_t* store_var;      // [ids][n_var]: heap position of the key's bytes of that column
  uint8_t* heap;
  uint32_t* new_len;        // [new slots + 1]: bytes a new key adds to the heap (all its variable-length columns), then its scan
  unsigned long long* var_bytes;   // [n_var] device totals of the bytes stored per column
};

// the reference's encoding of one row's key (hash_aggregate.cc:147-172): words[] must hold a.words zeros on entry
__device__ __forceinline__ void wide_encode(const WideArgs& a, int64_t row, uint64_t* words) {
  for (int k = 0; k < a.n_keys; ++k) {
    const WideKeyCol& c = a.cols[k];
    const bool valid = bitmap_get(c.vv, row);
    uint64_t v = 0;
    if (c.nbits == 0) {
      // [null byte][length]: the bytes are compared where they lie
      int pos = c.byte_off;
      if (!valid) words[pos >> 3] |= uint64_t(1) << ((pos & 7) * 8);
      ++pos;
      const uint32_t len = valid ? uint32_t(c.offsets[row + 1] - c.offsets[row]) : 0u;
      for (int b = 0; b < 4; ++b, ++pos) words[pos >> 3] |= uint64_t((len >> (8 * b)) & 0xFFu) << ((pos & 7) * 8);
      continue;
    }
    if (valid) {
      switch (c.nbits) {
        case 1: v = bitmap_get(c.bits, row); break;
        case 8: v = reinterpret_cast<const uint8_t*>(c.data)[row]; break;
        case 16: v = reinterpret_cast<const uint16_t*>(c.data)[row]; break;
        case 32: v = reinterpret_cast<const uint32_t*>(c.data)[row]; break;
        default: v = reinterpret_cast<const uint64_t*>(c.data)[row]; break;
      }
    }
    const int nbytes = c.nbits == 1 ? 1 : c.nbits / 8;
    // byte string: [null byte = 1 when null][value bytes, little endian]
    int pos = c.byte_off;
    if (!valid) words[pos >> 3] |= uint64_t(1) << ((pos & 7) * 8);
    ++pos;
    for (int b = 0; b < nbytes; ++b, ++pos) words[pos >> 3] |= ((v >> (8 * b)) & 0xFFull) << ((pos & 7) * 8);
  }
}

__device__ __forceinline__ uint64_t wide_hash(const uint64_t* words, int n) {
  uint64_t h = 0x9E3779B97F4A7C15ull;
  for (int i = 0; i < n; ++i) h = mix64(h ^ words[i]) + 0x9E3779B97F4A7C15ull * uint64_t(i + 1);
  return mix64(h);
}

// the bytes of a variable-length value folded into a hash, eight at a time
__device__ __forceinline__ uint64_t wide_hash_bytes(uint64_t h, const uint8_t* p, uint32_t len) {
  for (uint32_t i = 0; i < len; i += 8) {
    uint64_t w = 0;
    const uint32_t m = min(8u, len - i);
    for (uint32_t j = 0; j < m; ++j) w |= uint64_t(p[i + j]) << (8 * j);
    h = mix64(h ^ w) + 0x9E3779B97F4A7C15ull;
  }
  return h;
}
__device__ __forceinline__ bool wide_same_bytes(const uint8_t* x, const uint8_t* y, uint32_t len) {
  for (uint32_t i = 0; i < len; ++i)
    if (x[i] != y[i]) return false;
  return true;
}
// length field of a variable-length key column inside an encoded key
__device__ __forceinline__ uint32_t wide_len_of(const uint64_t* words, int byte_off) {
  uint32_t len = 0;
  int pos = byte_off + 1;
  for (int b = 0; b < 4; ++b, ++pos) len |= uint32_t((words[pos >> 3] >> ((pos & 7) * 8)) & 0xFFull) << (8 * b);
  return len;
}
// hash of a row's key: the encoded words, then the bytes of every variable-length column
__device__ __forceinline__ uint64_t wide_hash_row(const WideArgs& a, int64_t row, const uint64_t* words) {
  uint64_t h = wide_hash(words, a.words);
  for (int v = 0; v < a.n_var; ++v) {
    const WideKeyCol& c = a.cols[a.var_col[v]];
    const uint32_t len = wide_len_of(words, c.byte_off);
    if (len) h = mix64(wide_hash_bytes(h, c.bytes + c.offsets[row], len));
  }
  return h;
}

__device__ __forceinline__ uint64_t wide_ref(uint32_t kind, uint32_t tag, uint32_t payload) {
  return (uint64_t(kind) << 63) | (uint64_t(tag) << 32) | payload;
}
__device__ __forceinline__ uint32_t wide_tag(uint64_t h) {
  const uint32_t t = uint32_t(h >> 33);            // 31 bits
  return t == 0x7FFFFFFFu ? 0x7FFFFFFEu : t;       // {kind 1, tag all ones, payload all ones} would read as EMPTY
}

__device__ __forceinline__ bool wide_same_key(const WideArgs& a, uint64_t ref, const uint64_t* words, int64_t row) {
  if (ref >> 63) {
    const uint32_t id = uint32_t(ref);
    const uint64_t* k = a.store + uint64_t(id) * a.words;
    for (int i = 0; i < a.words; ++i)
      if (k[i] != words[i]) return false;
    for (int v = 0; v < a.n_var; ++v) {   // equal words: same null flags and lengths
      const WideKeyCol& c = a.cols[a.var_col[v]];
      const uint32_t len = wide_len_of(words, c.byte_off);
      if (len && !wide_same_bytes(c.bytes + c.offsets[row], a.heap + a.store_var[uint64_t(id) * a.n_var + v], len)) return false;
    }
    return true;
  }
  const int64_t q = int64_t(uint32_t(ref));
  uint64_t other[kWideWords];
  for (int i = 0; i < a.words; ++i) other[i] = 0;
  wide_encode(a, q, other);
  for (int i = 0; i < a.words; ++i)
    if (other[i] != words[i]) return false;
  for (int v = 0; v < a.n_var; ++v) {
    const WideKeyCol& c = a.cols[a.var_col[v]];
    const uint32_t len = wide_len_of(words, c.byte_off);
    if (len && !wide_same_bytes(c.bytes + c.offsets[row], c.bytes + c.offsets[q], len)) return false;
  }
  return true;
}

// pass 1: every row finds or claims the slot of its key
__global__ void wide_insert_kernel(const __grid_constant__ WideArgs a) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  if (a.select && !((a.select[row >> 5] >> (row & 31)) & 1u)) {
    a.row_slot[row] = 0xFFFFFFFFu;
    return;
  }
  uint64_t words[kWideWords];
  for (int i = 0; i < a.words; ++i) words[i] = 0;
  wide_encode(a, row, words);
  const uint64_t h = wide_hash_row(a, row, words);
  const uint32_t tag = wide_tag(h);
  uint64_t i = h & (a.cap - 1);
  for (int probe = 0; probe < kWideProbe; ++probe) {
    uint64_t cur = *reinterpret_cast<volatile uint64_t*>(a.slots + i);
    if (cur == kWideEmpty) {
      cur = atomicCAS(reinterpret_cast<unsigned long long*>(a.slots + i), (unsigned long long)kWideEmpty,
                      (unsigned long long)wide_ref(0, tag, uint32_t(row)));
      if (cur == kWideEmpty) {
        a.new_slots[atomicAdd(a.counters + 1, 1u)] = uint32_t(i);
        a.row_slot[row] = uint32_t(i);
        return;
      }
    }
    if (uint32_t(cur >> 32 & 0x7FFFFFFFu) == tag && wide_same_key(a, cur, words, row)) {
      a.row_slot[row] = uint32_t(i);
      return;
    }
    i = (i + 1) & (a.cap - 1);
  }
  atomicExch(a.counters + 2, 1u);   // cannot happen: the host keeps the load factor <= 1/2
}

// pass 2a (variable-length keys only): bytes every new key adds to the heap; per-column totals
__global__ void wide_new_len_kernel(const __grid_constant__ WideArgs a, uint32_t n_new) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_new) return;
  const int64_t row = int64_t(uint32_t(a.slots[a.new_slots[j]]));
  uint32_t total = 0;
  for (int v = 0; v < a.n_var; ++v) {
    const WideKeyCol& c = a.cols[a.var_col[v]];
    const uint32_t len = bitmap_get(c.vv, row) ? uint32_t(c.offsets[row + 1] - c.offsets[row]) : 0u;
    total += len;
    if (len) atomicAdd(a.var_bytes + v, (unsigned long long)len);
  }
  a.new_len[j] = total;
}

// pass 2: the slots claimed by this batch get their ids; their keys move into the store (and their bytes into the
// heap, at heap_base + the scanned new_len)
__global__ void wide_assign_kernel(const __grid_constant__ WideArgs a, uint32_t n_new, uint32_t id_base, uint64_t heap_base) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_new) return;
  const uint32_t slot = a.new_slots[j];
  const uint64_t ref = a.slots[slot];
  const int64_t row = int64_t(uint32_t(ref));
  uint64_t words[kWideWords];
  for (int i = 0; i < a.words; ++i) words[i] = 0;
  wide_encode(a, row, words);
  const uint32_t id = id_base + j;
  uint64_t* dst = a.store + uint64_t(id) * a.words;
  for (int i = 0; i < a.words; ++i) dst[i] = words[i];
  if (a.n_var) {
    uint64_t at = heap_base + a.new_len[j];
    for (int v = 0; v < a.n_var; ++v) {
      const WideKeyCol& c = a.cols[a.var_col[v]];
      const uint32_t len = wide_len_of(words, c.byte_off);
      a.store_var[uint64_t(id) * a.n_var + v] = at;
      const uint8_t* src = c.bytes + c.offsets[row];
      for (uint32_t b = 0; b < len; ++b) a.heap[at + b] = src[b];
      at += len;
    }
  }
  a.slots[slot] = wide_ref(1, uint32_t(ref >> 32) & 0x7FFFFFFFu, id);
}

// pass 3: row -> id
__global__ void wide_ids_kernel(const __grid_constant__ WideArgs a) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row >= a.n) return;
  const uint32_t slot = a.row_slot[row];
  a.ids[row] = slot == 0xFFFFFFFFu ? 0u : uint32_t(a.slots[slot]);
}

// a grown table: every stored key back into its slot
struct WideVarInfo {
  int32_t n_var;
  int32_t byte_off[kWideMaxKeys];   // of variable-length column v inside the encoded key
  const uint64_t* store_var;
  const uint8_t* heap;
};
__global__ void wide_rehash_kernel(const uint64_t* store, int words, uint32_t n_ids, uint64_t* slots, uint64_t cap, uint32_t* counters,
                                   const __grid_constant__ WideVarInfo vi) {
  const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= n_ids) return;
  uint64_t h = wide_hash(store + uint64_t(id) * words, words);
  for (int v = 0; v < vi.n_var; ++v) {   // == wide_hash_row
    const uint32_t len = wide_len_of(store + uint64_t(id) * words, vi.byte_off[v]);
    if (len) h = mix64(wide_hash_bytes(h, vi.heap + vi.store_var[uint64_t(id) * vi.n_var + v], len));
  }
  uint64_t i = h & (cap - 1);
  for (int probe = 0; probe < kWideProbe; ++probe) {
    if (atomicCAS(reinterpret_cast<unsigned long long*>(slots + i), (unsigned long long)kWideEmpty,
                  (unsigned long long)wide_ref(1, wide_tag(h), id)) == kWideEmpty)
      return;
    i = (i + 1) & (cap - 1);
  }
  atomicExch(counters + 2, 2u);
}

// finalize: ids of the groups (as emitted for the uint32 stand-in key) -> the key columns (KeyEncoder::Decode,
// hash_aggregate.cc:65-94, 174-192)
struct WideDecodeArgs {
  const uint32_t* ids;
  uint32_t n_groups;
  const uint64_t* store;
  int32_t words, n_keys;
  int32_t nbits[kWideMaxKeys], byte_off[kWideMaxKeys];
  void* data[kWideMaxKeys];           // variable-length key: the int32 offsets [n_groups + 1]
  uint32_t* validity[kWideMaxKeys];   // nullable
  uint32_t* null_counts;              // [kWideMaxKeys]
  // variable-length keys
  int32_t n_var;
  int32_t var_index[kWideMaxKeys];    // per key column, -1 for fixed-width
  uint8_t* out_bytes[kWideMaxKeys];
  const uint64_t* store_var;
  const uint8_t* heap;
};

__global__ void wide_decode_kernel(const __grid_constant__ WideDecodeArgs d) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in = g < d.n_groups;
  const int lane = threadIdx.x & 31;
  const uint64_t* words = d.store + uint64_t(in ? d.ids[g] : 0) * d.words;
  for (int k = 0; k < d.n_keys; ++k) {
    int pos = d.byte_off[k];
    bool valid = false;
    uint64_t v = 0;
    if (in) {
      valid = ((words[pos >> 3] >> ((pos & 7) * 8)) & 0xFFull) == 0;
      ++pos;
      const int nbytes = d.nbits[k] == 0 ? 4 : (d.nbits[k] == 1 ? 1 : d.nbits[k] / 8);   // variable-length: the length field
      for (int b = 0; b < nbytes; ++b, ++pos) v |= ((words[pos >> 3] >> ((pos & 7) * 8)) & 0xFFull) << (8 * b);
    }
    if (d.nbits[k] == 1) {
      const unsigned word = __ballot_sync(0xffffffffu, in && valid && v);
      if (lane == 0 && in) reinterpret_cast<uint32_t*>(d.data[k])[g >> 5] = word;
    } else if (in) {
      switch (d.nbits[k]) {
        case 0: reinterpret_cast<uint32_t*>(d.data[k])[g] = valid ? uint32_t(v) : 0u; break;   // lengths: scanned into offsets next
        case 8: reinterpret_cast<uint8_t*>(d.data[k])[g] = uint8_t(v); break;
        case 16: reinterpret_cast<uint16_t*>(d.data[k])[g] = uint16_t(v); break;
        case 32: reinterpret_cast<uint32_t*>(d.data[k])[g] = uint32_t(v); break;
        default: reinterpret_cast<uint64_t*>(d.data[k])[g] = v; break;
      }
    }
    const unsigned vw = __ballot_sync(0xffffffffu, in && valid);
    const unsigned inw = __ballot_sync(0xffffffffu, in);
    if (lane == 0 && inw) {
      if (d.validity[k]) d.validity[k][g >> 5] = vw;
      const int nulls = __popc(inw & ~vw);
      if (nulls) atomicAdd(d.null_counts + k, uint32_t(nulls));
    }
  }
}

// variable-length key columns, after the lengths written by wide_decode_kernel were scanned into offsets: the bytes
__global__ void wide_decode_bytes_kernel(const __grid_constant__ WideDecodeArgs d) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= d.n_groups) return;
  const uint32_t id = d.ids[g];
  for (int k = 0; k < d.n_keys; ++k) {
    if (d.nbits[k] != 0) continue;
    const int32_t* off = reinterpret_cast<const int32_t*>(d.data[k]);
    const uint32_t len = uint32_t(off[g + 1] - off[g]);
    const uint8_t* src = d.heap + d.store_var[uint64_t(id) * d.n_var + d.var_index[k]];
    uint8_t* dst = d.out_bytes[k] + off[g];
    for (uint32_t b = 0; b < len; ++b) dst[b] = src[b];
  }
}
