// jpeg.cu -- kernel (d): JPEG colour transform, FDCT, quantisation, Huffman coding.
#include "../../include/stitch_b200.h"
#include "kernels.h"

int stitch_jpeg_run_band(stitch_job *, uint32_t, uint32_t, uint8_t *, size_t, size_t *, stitch_timings *) {
  return STITCH_E_STATE;
}
extern "C" int stitch_jpeg_header(uint32_t, uint32_t, uint32_t, uint8_t *, size_t, size_t *) { return STITCH_E_STATE; }
