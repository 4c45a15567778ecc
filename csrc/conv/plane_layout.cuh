// "Plane layout": the activation format that makes a 5x5 SAME convolution a zero-copy tensor-core operand.
//
// Per image, C channels are split into C/8 planes.  A plane is a 16 x 16 grid of pixels, 8 bf16 channels
// (16 bytes) per pixel, 4096 bytes per plane.  Plane row r stores image row r-2 (rows 0 and 1 are zero), plane
// column c stores image column c (columns 14 and 15 are zero).  Consequences, for a 14 x 14 image:
//
//   * a pixel's 16-byte chunk is exactly one row of a UMMA "core matrix" (8 rows x 16 bytes, contiguous), so
//     a plane *is* a canonical no-swizzle K-major operand whose M rows are pixels (SBO = 128 B) and whose
//     K chunks are planes (LBO = 4096 B) — or, transposed, a canonical MN-major operand whose K is pixels;
//   * output pixel m = h*16 + w under filter tap (kh, kw) reads plane pixel m + kh*16 + kw - 2: a constant
//     offset.  The tap is applied by re-basing the descriptor start address; nothing is gathered or copied;
//   * the zero columns 14-15 serve as right halo of one row and left halo of the next, the zero rows 0-1 of
//     the following plane serve as bottom halo, so SAME padding costs no extra memory traffic either.
//
// The whole image is one contiguous block (16 KiB for 32 channels, 32 KiB for 64), fetched with a single
// cp.async.bulk.  Verified on silicon by tools/umma_probe.py (conv_shift, mnmajor_noswz_*).
#pragma once
#include <stdint.h>

namespace dtm {

constexpr int PL_GRID = 16;                    // plane is 16 x 16 pixels
constexpr int PL_PIXELS = PL_GRID * PL_GRID;   // 256
constexpr int PL_PLANE_BYTES = PL_PIXELS * 16; // 4096
constexpr int PL_ROW_OFF = 2;                  // image row h lives in plane row h + 2
constexpr int IMG = 14;                        // spatial size of both tensor-core conv operands (14 x 14)

__host__ __device__ constexpr int plane_image_bytes(int channels) { return (channels / 8) * PL_PLANE_BYTES; }

// Byte offset of channel c of image pixel (h, w) inside one image of a plane-layout tensor.
__host__ __device__ constexpr int plane_offset(int h, int w, int c) {
  return (c >> 3) * PL_PLANE_BYTES + ((h + PL_ROW_OFF) * PL_GRID + w) * 16 + (c & 7) * 2;
}

// Zero guard kept in shared memory before and after the image buffers: taps reach at most 2 rows + 2 pixels
// (34 pixels = 544 B) before the first plane and 66 pixels (1056 B) past the last one.
constexpr int PL_GUARD_BYTES = 2048;

}  // namespace dtm
