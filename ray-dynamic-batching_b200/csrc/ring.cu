// Request ring: the per-model FIFO of the scheduler (293-project/src/scheduler.py:190-372 RequestQueue over a Ray queue
// actor) as a lock-free ring in caller-provided memory -- normally a shared-memory segment that the submitting process
// and every model-actor process map, and that each actor registers with CUDA so a batch goes to its GPU by DMA straight
// from where the request was written.
//
// What the reference does per request: three blocking RPCs and one pickled 602 KB tensor (scheduler.py:238-254 through
// python/ray/util/queue.py:80-109); per batch: one RPC per item plus a Python loop (scheduler.py:274-289).  Here submit
// is one memcpy of the request's bytes into the next slot plus a 128-byte descriptor, and batch formation is one call
// that claims a run of consecutive slots, applies the reference's stale-request rule (scheduler.py:281) and hands back
// slot numbers -- no per-request work on the Python side in either direction.
//
// Protocol (bounded MPMC queue with per-slot sequence numbers): slot s = seq % nslots carries `state`:
//   state == seq            free for the producer of `seq`
//   state == seq + 1        published (payload and descriptor complete)
//   state == seq + nslots   released by the consumer = free for the next lap
// A producer claims `head` by CAS after seeing its slot free, fills it, publishes.  A consumer claims [tail, tail + k) by
// CAS over a prefix of published slots, and OWNS those slots until it releases them -- after its host-to-device copy --
// so a slow actor can never have its requests overwritten: a producer that meets an unreleased slot reports "full",
// which the scheduler counts as a dropped request exactly like the reference's full queue (scheduler.py:241-243).
//
// Host code only (no kernels); compiled into libb200infer.so because it is the other half of the same boundary.
#include <atomic>
#if defined(__x86_64__) || defined(_M_X64)
#include <emmintrin.h>
#endif
#include <algorithm>
#include <math.h>
#include <new>
#include <vector>

#include "common.h"

namespace {

using b200::fail;

constexpr uint64_t RING_MAGIC = 0x42323030524e4731ull;  // "B200RNG1"
constexpr int LAT_CAP = 1024;                            // latency ring (the reference keeps 900-1000, scheduler.py:333-336)

struct alignas(128) RingHeader {
  uint64_t magic;
  uint32_t nslots, slot_bytes, max_queued, reserved;
  uint64_t desc_off, payload_off, total_bytes;
  alignas(128) std::atomic<uint64_t> head;   // next sequence number a producer will claim
  alignas(128) std::atomic<uint64_t> tail;   // next sequence number a consumer will claim
  alignas(128) std::atomic<uint64_t> total_requests;
  std::atomic<uint64_t> dropped_requests;
  std::atomic<uint64_t> discarded_stale;
  std::atomic<uint64_t> slo_violations;
  std::atomic<uint64_t> completed_requests;
  std::atomic<uint64_t> completed_batches;
  std::atomic<uint64_t> lat_count;           // latencies ever appended; entry i lives at lat[i % LAT_CAP]
  std::atomic<uint64_t> last_latency_bits;   // double, bit-cast
  std::atomic<uint64_t> max_latency_bits;    // largest batch latency ever recorded (non-negative doubles order like their bits)
  double lat[LAT_CAP];
};

struct alignas(128) RingDesc {
  std::atomic<uint64_t> state;
  double arrival;            // time.time() of the submit (scheduler.py:246)
  uint32_t nbytes;
  uint16_t h, w;
  uint8_t c, pad[3];
  char id[B200_RING_ID_BYTES];
};
static_assert(sizeof(RingDesc) == 128, "descriptor is one 128-byte line");
static_assert(std::atomic<uint64_t>::is_always_lock_free, "shared-memory atomics must be lock-free");

inline RingHeader* hdr_of(void* mem) { return static_cast<RingHeader*>(mem); }
inline RingDesc* descs_of(RingHeader* h) { return reinterpret_cast<RingDesc*>(reinterpret_cast<char*>(h) + h->desc_off); }
inline char* payload_of(RingHeader* h) { return reinterpret_cast<char*>(h) + h->payload_off; }

inline int check(void* mem, RingHeader** out) {
  if (!mem) return fail(B200_ERR_INVALID, "ring: null memory");
  RingHeader* h = hdr_of(mem);
  if (h->magic != RING_MAGIC) return fail(B200_ERR_INVALID, "ring: memory does not hold an initialised ring");
  *out = h;
  return B200_OK;
}

inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// A request's bytes are written once by a client and read once by a GPU's copy engine; nothing on the host reads them in
// between.  Non-temporal stores skip the read-for-ownership of every destination line and leave the payload in DRAM, where
// the DMA finds it, instead of dirty in the writer's caches: with twenty clients feeding eight GPUs the host's memory
// traffic per request drops by a third.  dst is 256-byte aligned (slots are), src is whatever the caller has.
inline void stream_copy(void* dst, const void* src, size_t n) {
#if defined(__x86_64__) || defined(_M_X64)
  if (n >= 4096 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    const size_t blocks = n / 64;
    __m128i* d = static_cast<__m128i*>(dst);
    const __m128i* s = static_cast<const __m128i*>(src);
    for (size_t i = 0; i < blocks; ++i) {
      const __m128i a = _mm_loadu_si128(s + 4 * i), b = _mm_loadu_si128(s + 4 * i + 1);
      const __m128i c = _mm_loadu_si128(s + 4 * i + 2), e = _mm_loadu_si128(s + 4 * i + 3);
      _mm_stream_si128(d + 4 * i, a);
      _mm_stream_si128(d + 4 * i + 1, b);
      _mm_stream_si128(d + 4 * i + 2, c);
      _mm_stream_si128(d + 4 * i + 3, e);
    }
    const size_t done = blocks * 64;
    if (done < n) memcpy(static_cast<char*>(dst) + done, static_cast<const char*>(src) + done, n - done);
    _mm_sfence();                      // the streamed lines are globally visible before the slot is published
    return;
  }
#endif
  memcpy(dst, src, n);
}

}  // namespace

extern "C" {

size_t b200_ring_bytes(uint32_t nslots, uint32_t slot_bytes) {
  const size_t slot = round_up(slot_bytes, 256);
  return round_up(sizeof(RingHeader), 4096) + round_up(size_t(nslots) * sizeof(RingDesc), 4096) + size_t(nslots) * slot;
}

int b200_ring_init(void* mem, size_t bytes, uint32_t nslots, uint32_t slot_bytes, uint32_t max_queued) {
  if (!mem || nslots == 0 || slot_bytes == 0) return fail(B200_ERR_INVALID, "ring_init: bad arguments");
  if (max_queued == 0 || max_queued > nslots) return fail(B200_ERR_INVALID, "ring_init: need 0 < max_queued <= nslots");
  if ((reinterpret_cast<uintptr_t>(mem) & 127) != 0) return fail(B200_ERR_INVALID, "ring_init: memory must be 128-byte aligned");
  if (bytes < b200_ring_bytes(nslots, slot_bytes)) return fail(B200_ERR_INVALID, "ring_init: memory too small (b200_ring_bytes)");
  RingHeader* h = new (mem) RingHeader;
  h->nslots = nslots;
  h->slot_bytes = static_cast<uint32_t>(round_up(slot_bytes, 256));
  h->max_queued = max_queued;
  h->reserved = 0;
  h->desc_off = round_up(sizeof(RingHeader), 4096);
  h->payload_off = h->desc_off + round_up(size_t(nslots) * sizeof(RingDesc), 4096);
  h->total_bytes = b200_ring_bytes(nslots, slot_bytes);
  h->head.store(0);
  h->tail.store(0);
  h->total_requests.store(0);
  h->dropped_requests.store(0);
  h->discarded_stale.store(0);
  h->slo_violations.store(0);
  h->completed_requests.store(0);
  h->completed_batches.store(0);
  h->lat_count.store(0);
  h->last_latency_bits.store(0);
  h->max_latency_bits.store(0);
  RingDesc* d = descs_of(h);
  for (uint32_t i = 0; i < nslots; ++i) {
    new (&d[i]) RingDesc;
    d[i].state.store(i, std::memory_order_relaxed);
  }
  std::atomic_thread_fence(std::memory_order_seq_cst);
  h->magic = RING_MAGIC;
  return B200_OK;
}

int b200_ring_layout(void* mem, uint32_t* nslots, uint32_t* slot_bytes, uint32_t* max_queued, uint64_t* payload_off,
                     uint64_t* total_bytes) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  if (nslots) *nslots = h->nslots;
  if (slot_bytes) *slot_bytes = h->slot_bytes;
  if (max_queued) *max_queued = h->max_queued;
  if (payload_off) *payload_off = h->payload_off;
  if (total_bytes) *total_bytes = h->total_bytes;
  return B200_OK;
}

long long b200_ring_qsize(void* mem) {
  RingHeader* h;
  if (check(mem, &h) != 0) return -1;
  const uint64_t t = h->tail.load(std::memory_order_acquire);
  const uint64_t hd = h->head.load(std::memory_order_acquire);
  return hd >= t ? static_cast<long long>(hd - t) : 0;
}

// Returns 1 when the request was queued, 0 when the queue is full (request dropped, counted), negative on misuse.
int b200_ring_push(void* mem, const char* request_id, const void* payload, uint32_t nbytes, int h_px, int w_px, int c,
                   double arrival) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  if (!payload || nbytes == 0 || nbytes > h->slot_bytes)
    return fail(B200_ERR_INVALID, "ring_push: request of " + std::to_string(nbytes) + " bytes does not fit a slot of " +
                                      std::to_string(h->slot_bytes));
  const size_t idlen = request_id ? strlen(request_id) : 0;
  if (idlen >= B200_RING_ID_BYTES) return fail(B200_ERR_INVALID, "ring_push: request id longer than B200_RING_ID_BYTES - 1");
  RingDesc* descs = descs_of(h);
  uint64_t pos = h->head.load(std::memory_order_relaxed);
  for (;;) {
    const uint64_t t = h->tail.load(std::memory_order_acquire);
    if (pos >= t && pos - t >= h->max_queued) {
      h->dropped_requests.fetch_add(1, std::memory_order_relaxed);
      return 0;                                     // scheduler.py:241-243: queue full, drop
    }
    RingDesc& d = descs[pos % h->nslots];
    const uint64_t s = d.state.load(std::memory_order_acquire);
    if (s == pos) {
      if (h->head.compare_exchange_weak(pos, pos + 1, std::memory_order_acq_rel, std::memory_order_relaxed)) break;
    } else if (s < pos) {                           // slot of the previous lap not released yet: its actor still owns it
      h->dropped_requests.fetch_add(1, std::memory_order_relaxed);
      return 0;
    } else {
      pos = h->head.load(std::memory_order_relaxed);  // another producer took it
    }
  }
  RingDesc& d = descs[pos % h->nslots];
  stream_copy(payload_of(h) + size_t(pos % h->nslots) * h->slot_bytes, payload, nbytes);
  d.arrival = arrival;
  d.nbytes = nbytes;
  d.h = static_cast<uint16_t>(h_px);
  d.w = static_cast<uint16_t>(w_px);
  d.c = static_cast<uint8_t>(c);
  memcpy(d.id, request_id ? request_id : "", idlen);
  d.id[idlen] = 0;
  d.state.store(pos + 1, std::memory_order_release);
  h->total_requests.fetch_add(1, std::memory_order_relaxed);
  return 1;
}

// n requests of equal size laid out `stride` bytes apart; ids is n entries of B200_RING_ID_BYTES (NUL-terminated), or NULL
// (ids "<prefix><first_index + i>").  Returns how many were queued (a full queue stops the run), negative on misuse.
int b200_ring_push_many(void* mem, int n, const char* ids, const char* id_prefix, long long first_index,
                        const void* payloads, size_t stride, uint32_t nbytes, int h_px, int w_px, int c, double arrival) {
  int done = 0;
  char idbuf[B200_RING_ID_BYTES];
  for (int i = 0; i < n; ++i) {
    const char* id;
    if (ids) {
      id = ids + size_t(i) * B200_RING_ID_BYTES;
    } else {
      snprintf(idbuf, sizeof(idbuf), "%s%lld", id_prefix ? id_prefix : "r", first_index + i);
      id = idbuf;
    }
    const int rc = b200_ring_push(mem, id, static_cast<const char*>(payloads) + size_t(i) * stride, nbytes, h_px, w_px, c,
                                  arrival);
    if (rc < 0) return rc;
    if (rc == 0) {                                  // full: the rest of the burst is dropped with it (one was counted by push)
      RingHeader* h = hdr_of(mem);
      h->dropped_requests.fetch_add(uint64_t(n - i - 1), std::memory_order_relaxed);
      break;
    }
    ++done;
  }
  return done;
}

// Claims up to max_items requests from the front of the queue (FIFO).  A claimed request whose deadline cannot be met,
//   arrival + slo_s < now + expected_latency_s      (scheduler.py:281)
// is released at once and counted in *discarded; the others are returned: their slot numbers (-> payload address), arrival
// times, sizes and ids.  The caller owns the returned slots until b200_ring_release.  Returns the number returned.
int b200_ring_pop(void* mem, int max_items, double now, double expected_latency_s, double slo_s, uint32_t* slots,
                  uint64_t* seqs, double* arrivals, int32_t* hw, uint32_t* nbytes, char* ids, int* discarded) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  if (discarded) *discarded = 0;
  if (max_items <= 0) return 0;
  if (!slots || !seqs) return fail(B200_ERR_INVALID, "ring_pop: null output");
  RingDesc* descs = descs_of(h);
  uint64_t t = h->tail.load(std::memory_order_relaxed);
  uint64_t k = 0;
  for (;;) {
    const uint64_t hd = h->head.load(std::memory_order_acquire);
    const uint64_t avail = hd > t ? std::min<uint64_t>(hd - t, uint64_t(max_items)) : 0;
    k = 0;
    while (k < avail && descs[(t + k) % h->nslots].state.load(std::memory_order_acquire) == t + k + 1) ++k;
    if (k == 0) return 0;                            // empty, or the front request is still being written
    if (h->tail.compare_exchange_weak(t, t + k, std::memory_order_acq_rel, std::memory_order_relaxed)) break;
  }
  int out = 0, dropped = 0;
  for (uint64_t i = 0; i < k; ++i) {
    const uint64_t seq = t + i;
    RingDesc& d = descs[seq % h->nslots];
    if (d.arrival + slo_s < now + expected_latency_s) {
      d.state.store(seq + h->nslots, std::memory_order_release);
      ++dropped;
      continue;
    }
    slots[out] = static_cast<uint32_t>(seq % h->nslots);
    seqs[out] = seq;
    if (arrivals) arrivals[out] = d.arrival;
    if (hw) { hw[2 * out] = d.h; hw[2 * out + 1] = d.c == 1 ? -static_cast<int32_t>(d.w) : d.w; }   // grayscale: negative width (b200_collate)
    if (nbytes) nbytes[out] = d.nbytes;
    if (ids) memcpy(ids + size_t(out) * B200_RING_ID_BYTES, d.id, B200_RING_ID_BYTES);
    ++out;
  }
  if (dropped) h->discarded_stale.fetch_add(dropped, std::memory_order_relaxed);
  if (discarded) *discarded = dropped;
  return out;
}

int b200_ring_release(void* mem, const uint64_t* seqs, int n) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  RingDesc* descs = descs_of(h);
  for (int i = 0; i < n; ++i) descs[seqs[i] % h->nslots].state.store(seqs[i] + h->nslots, std::memory_order_release);
  return B200_OK;
}

// record_batch_completion (scheduler.py:324-341): one latency entry per request of the batch, violations counted against
// the SLO.  Returns the number of violations.
int b200_ring_record(void* mem, double latency_ms, int n_requests, double slo_ms) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  if (n_requests < 0) return fail(B200_ERR_INVALID, "ring_record: negative request count");
  uint64_t bits;
  memcpy(&bits, &latency_ms, 8);
  h->last_latency_bits.store(bits, std::memory_order_relaxed);
  if (latency_ms > 0) {
    uint64_t cur = h->max_latency_bits.load(std::memory_order_relaxed);
    while (bits > cur && !h->max_latency_bits.compare_exchange_weak(cur, bits, std::memory_order_relaxed)) {}
  }
  const uint64_t first = h->lat_count.fetch_add(uint64_t(n_requests), std::memory_order_acq_rel);
  for (int i = 0; i < n_requests; ++i) h->lat[(first + i) % LAT_CAP] = latency_ms;
  int viol = 0;
  if (latency_ms > slo_ms) {
    viol = n_requests;
    h->slo_violations.fetch_add(uint64_t(n_requests), std::memory_order_relaxed);
  }
  h->completed_requests.fetch_add(uint64_t(n_requests), std::memory_order_relaxed);
  h->completed_batches.fetch_add(1, std::memory_order_relaxed);
  return viol;
}

// The latency window get_stats reads, oldest first (what the reference's `metrics['latencies']` list holds): up to `cap`
// values into `out`; returns how many.  Lets a caller that shards one model's queue over several rings merge their windows.
int b200_ring_latencies(void* mem, double* out, int cap) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  const uint64_t n = h->lat_count.load(std::memory_order_acquire);
  uint64_t win = n <= 1000 ? n : 901 + (n - 1001) % 100;
  if (win > uint64_t(cap)) win = uint64_t(cap);
  for (uint64_t i = 0; i < win; ++i) out[i] = h->lat[(n - win + i) % LAT_CAP];
  return static_cast<int>(win);
}

// get_stats (scheduler.py:343-372).  out[0..10] = total_requests, dropped_requests, slo_violations, queue_size,
// last_latency, avg_latency, p95_latency, p99_latency, discarded_stale, completed_requests, completed_batches.
// The window is what the reference's list holds after the same number of appends: all of them up to 1000, then the last
// 901..1000 (it drops the oldest 100 whenever it reaches 1000, scheduler.py:333-334); percentiles index the sorted
// window at int(n * 0.95) / int(n * 0.99).
int b200_ring_stats(void* mem, double* out, int n_out) {
  RingHeader* h;
  B200_TRY(check(mem, &h));
  if (!out || n_out < 11) return fail(B200_ERR_INVALID, "ring_stats: need room for 11 values");
  out[0] = double(h->total_requests.load());
  out[1] = double(h->dropped_requests.load());
  out[2] = double(h->slo_violations.load());
  out[3] = double(b200_ring_qsize(mem));
  const uint64_t bits = h->last_latency_bits.load();
  memcpy(&out[4], &bits, 8);
  const uint64_t n = h->lat_count.load(std::memory_order_acquire);
  const uint64_t win = n <= 1000 ? n : 901 + (n - 1001) % 100;
  out[5] = out[6] = out[7] = 0.0;
  if (win > 0) {
    std::vector<double> v(win);
    // sum(latencies) / len(latencies) as CPython >= 3.12 evaluates it: compensated (Neumaier) summation in list order
    volatile double total = 0.0, comp = 0.0;
    for (uint64_t i = 0; i < win; ++i) {
      const double x = h->lat[(n - win + i) % LAT_CAP];
      v[i] = x;
      const double t = total + x;
      if (fabs(total) >= fabs(x)) comp = comp + ((total - t) + x);
      else comp = comp + ((x - t) + total);
      total = t;
    }
    out[5] = (total + comp) / double(win);
    std::sort(v.begin(), v.end());
    out[6] = v[size_t(double(win) * 0.95)];
    out[7] = v[size_t(double(win) * 0.99)];
  }
  out[8] = double(h->discarded_stale.load());
  out[9] = double(h->completed_requests.load());
  out[10] = double(h->completed_batches.load());
  if (n_out >= 12) {
    const uint64_t mb = h->max_latency_bits.load();
    memcpy(&out[11], &mb, 8);
  }
  return B200_OK;
}

}  // extern "C"
