// tf.train.shuffle_batch of cifar10cnn.py:82-90 (capacity 5384, min_after_dequeue 5000, one reader thread over
// the file list of :78-80) as one kernel over a record set that is resident in HBM.
//
// Law implemented (the queue with its reader always ahead, i.e. full at every dequeue):
//   * the reader walks the files in the epoch's file order (string_input_producer shuffles the file names once
//     per epoch) and inside a file record by record: stream position p -> record id
//         file_order[(p / per_file) % n_order] * per_file + p % per_file;
//   * the queue holds `capacity` record ids; one dequeue_many(batch) removes `batch` of them, each drawn uniformly
//     from what is left (a partial Fisher-Yates shuffle of the slot array: draw i swaps slot i with slot
//     i + floor(u_i * (capacity - i) / 2^32), u_i the top 32 bits of r_i), and the reader refills the freed slots
//     with the next `batch` stream positions.
//   Which slot a record sits in does not matter to a uniform draw, so refilling in place keeps the law.
// r_i = splitmix64(seed, step, i): a counter-based generator, so the host-side model in the tests reproduces every
// batch bit for bit and a restarted job can re-enter the sequence at any step.
//
// grid = batch CTAs: every CTA repeats the (tiny) draw on its own copy of the slot array in shared memory and then
// copies ONE record -- 3072 image bytes + its label -- so the 393 KB gather runs on 128 SMs at once; CTA 0 alone
// writes the new slot array into the OTHER buffer (slots_out != slots_in: no CTA may see a half-updated array).
#include "launch.h"

namespace b200 {
namespace {

__host__ __device__ inline uint64_t splitmix64(uint64_t seed, uint64_t step, uint64_t i) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (step * 65536ull + i + 1ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

struct ShuffleArgs {
  const uint8_t* images;      // [n_records][image_bytes]
  const int32_t* labels;      // [n_records]
  const int32_t* file_order;  // [n_order] file index per (epoch, position in the epoch's file list), flattened
  const int32_t* slots_in;    // [capacity] record ids in the queue
  int32_t* slots_out;         // [capacity] the queue after this dequeue + refill
  uint8_t* images_out;        // [batch][image_bytes]
  int32_t* labels_out;        // [batch]
  int64_t stream_pos;         // position of the next record the reader will enqueue
  uint64_t seed, step;
  int32_t n_order, per_file, capacity, batch, image_bytes;
};

__global__ void __launch_bounds__(256) shuffle_batch_kernel(ShuffleArgs a) {
  extern __shared__ int32_t slots[];
  __shared__ int32_t pick;
  const int tid = threadIdx.x, b = blockIdx.x;
  for (int i = tid; i < a.capacity; i += blockDim.x) slots[i] = a.slots_in[i];
  __syncthreads();
  if (tid == 0) {
    // draws 0..b decide which record this CTA emits; CTA 0 needs all of them for the write-back
    const int last = b == 0 ? a.batch - 1 : b;
    for (int i = 0; i <= last; ++i) {
      const uint32_t u = (uint32_t)(splitmix64(a.seed, a.step, (uint64_t)i) >> 32);
      const int j = i + (int)__umulhi(u, (uint32_t)(a.capacity - i));
      const int32_t t = slots[i]; slots[i] = slots[j]; slots[j] = t;
    }
    pick = slots[b];
  }
  __syncthreads();
  const int32_t rec = pick;
  const uint4* src = reinterpret_cast<const uint4*>(a.images + (size_t)rec * a.image_bytes);
  uint4* dst = reinterpret_cast<uint4*>(a.images_out + (size_t)b * a.image_bytes);
  for (int i = tid; i < a.image_bytes / 16; i += blockDim.x) dst[i] = __ldg(src + i);
  if (tid == 0) a.labels_out[b] = __ldg(a.labels + rec);
  if (b == 0) {
    // the reader refills the freed slots (0 .. batch-1 after the partial shuffle) with the next stream positions
    for (int i = tid; i < a.capacity; i += blockDim.x) {
      int32_t v = slots[i];
      if (i < a.batch) {
        const int64_t p = a.stream_pos + i;
        v = a.file_order[(p / a.per_file) % a.n_order] * a.per_file + (int32_t)(p % a.per_file);
      }
      a.slots_out[i] = v;
    }
  }
}
}  // namespace

int shuffle_batch(const uint8_t* images, const int32_t* labels, int64_t n_records, const int32_t* file_order,
                  int n_order, int per_file, const int32_t* slots_in, int32_t* slots_out, int capacity, uint64_t seed,
                  int64_t step, int64_t stream_pos, uint8_t* images_out, int32_t* labels_out, int batch, int image_bytes,
                  cudaStream_t st) {
  B200_REQUIRE(images && labels && file_order && slots_in && slots_out && images_out && labels_out,
               "shuffle_batch: null pointer");
  B200_REQUIRE(slots_in != slots_out, "shuffle_batch: slots_in and slots_out must be different buffers");
  B200_REQUIRE(batch > 0 && capacity > batch && capacity <= 12000, "shuffle_batch: need 0 < batch < capacity <= 12000");
  B200_REQUIRE(n_order > 0 && per_file > 0 && n_records > 0 && n_records % per_file == 0,
               "shuffle_batch: %lld records do not split into files of %d", (long long)n_records, per_file);
  B200_REQUIRE(image_bytes > 0 && image_bytes % 16 == 0 && ((uintptr_t)images & 15) == 0 && ((uintptr_t)images_out & 15) == 0,
               "shuffle_batch: records must be 16-byte aligned multiples of 16 bytes");
  B200_REQUIRE(step >= 0 && stream_pos >= 0, "shuffle_batch: negative step / stream position");
  ShuffleArgs a{images, labels, file_order, slots_in, slots_out, images_out, labels_out, stream_pos, seed,
                (uint64_t)step, n_order, per_file, capacity, batch, image_bytes};
  static PerDeviceOnce attr;
  if (attr.first())
    B200_CUDA_TRY(cudaFuncSetAttribute(shuffle_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024));
  shuffle_batch_kernel<<<batch, 256, (size_t)capacity * 4, st>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200
