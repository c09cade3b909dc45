// D2H of many regularly spaced pieces: one 2D copy per utterance (rows = the segments of the utterance, one time
// slice of each) against a loop of plain copies per piece against one big copy. Sizes of the bench batch.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/piece_copy_probe tools/piece_copy_probe.cu
#include <chrono>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
  const size_t utt = 1024, rows = 7, frame = 240 * 8;            // utterances, segments per utterance, bytes per frame
  const size_t seg_frames = 146, pitch = seg_frames * frame;      // one segment
  const size_t total = utt * rows * pitch;                        // ~2 GB
  char *d, *h;
  cudaMalloc(&d, total);
  cudaMallocHost(&h, total);
  cudaMemset(d, 1, total);
  cudaStream_t st;
  cudaStreamCreate(&st);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int K : {1, 2, 3, 4, 6}) {
    const size_t width = seg_frames / K * frame;  // one slice of every segment
    for (int mode = 0; mode < 3; ++mode) {
      float ms = 0; double api = 0; size_t calls = 0;
      for (int rep = 0; rep < 3; ++rep) {
        cudaStreamSynchronize(st);
        const double t0 = now();
        cudaEventRecord(a, st);
        cudaError_t e = cudaSuccess;
        calls = 0;
        if (mode == 0) {  // one 2D copy per utterance
          for (size_t u = 0; u < utt && e == cudaSuccess; ++u, ++calls)
            e = cudaMemcpy2DAsync(h + u * rows * pitch, pitch, d + u * rows * pitch, pitch, width, rows, cudaMemcpyDeviceToHost, st);
        } else if (mode == 1) {  // one copy per piece
          for (size_t i = 0; i < utt * rows && e == cudaSuccess; ++i, ++calls)
            e = cudaMemcpyAsync(h + i * pitch, d + i * pitch, width, cudaMemcpyDeviceToHost, st);
        } else {  // the same bytes as one contiguous copy
          e = cudaMemcpyAsync(h, d, width * utt * rows, cudaMemcpyDeviceToHost, st);
          calls = 1;
        }
        cudaEventRecord(b, st);
        api = now() - t0;
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&ms, a, b);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); break; }
      }
      printf("K=%d %-22s %6zu calls x %7zu B rows: enqueue %.3f ms, copy %.3f ms, %.1f GB/s\n", K,
             mode == 0 ? "2D copy per utterance" : mode == 1 ? "copy per piece" : "one contiguous copy", calls, width, api, ms,
             (double)width * utt * rows / ms * 1e-6);
    }
  }
  return 0;
}
