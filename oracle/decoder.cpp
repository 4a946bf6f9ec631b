// oracle/decoder.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// Velodyne HDL/VLP packet decode restated: what the reference's prebuilt lib/libPacketDecoder.so
// (PacketDecoder::DecodePacket / ProcessHDLPacket / PushFiringData / SplitFrame, declared in
// include/PacketDecoder.h:79-121, driven at src/pcapReader.cpp:79-102) does with 16db.xml.  Its source is
// not in the reference tree; this restatement was fitted to the binary itself and is PINNED: the golden
// fixtures under tests/golden/decoder_*.npz were produced by running that binary here
// (oracle/ref_decoder/), and tests/test_golden.py holds this function to them bit for bit (doubles).
//   tables: cos/sin((i / 100.0) * M_PI / 180.0), i = 0..36000  (HDL_Grabber_toRadians, PacketDecoder.h:15)
//   per return with distance != 0:  d = raw * 0.002;  xy = d * cos(v);  x = xy * sin(a);  y = xy * cos(a);
//   z = d * sin(v),  v = vertCorrection[laser id] in degrees, laser id = return slot (+32 for a 0xDDFF block);
//   ms = the packet's timestamp word; a block whose azimuth is below the previous block's starts a new frame.
#include <cmath>
#include <cstring>

#include "pcp_oracle.h"

extern "C" int64_t orc_decode_packets(const uint8_t* packets, int64_t n_packets, const float* vert_deg /*64*/, double* x, double* y,
                                      double* z, double* distance, uint8_t* intensity, uint8_t* laser_id, uint16_t* azimuth, uint32_t* ms,
                                      int64_t* frame_offsets, int64_t max_frames, int64_t* n_frames) {
  static double* cos_t = nullptr;
  static double* sin_t = nullptr;
  if (!cos_t) {
    cos_t = new double[36001];
    sin_t = new double[36001];
    for (int i = 0; i < 36001; ++i) {
      const double rad = ((double)i / 100.0) * M_PI / 180.0;
      cos_t[i] = cos(rad);
      sin_t[i] = sin(rad);
    }
  }
  double cv[64], sv[64];
  for (int l = 0; l < 64; ++l) {
    const double rad = ((double)vert_deg[l]) * M_PI / 180.0;
    cv[l] = cos(rad);
    sv[l] = sin(rad);
  }
  int64_t n = 0, nf = 0;
  unsigned last_az = 0;
  if (max_frames > 0) frame_offsets[0] = 0;
  for (int64_t p = 0; p < n_packets; ++p) {
    const uint8_t* pk = packets + p * 1206;
    uint32_t ts;
    memcpy(&ts, pk + 1200, 4);
    for (int b = 0; b < 12; ++b) {
      const uint8_t* blk = pk + b * 100;
      uint16_t flag, az;
      memcpy(&flag, blk, 2);
      memcpy(&az, blk + 2, 2);
      if ((unsigned)az < last_az) {  // SplitFrame
        ++nf;
        if (nf < max_frames) frame_offsets[nf] = n;
      }
      last_az = az;
      const int base = flag == 0xDDFF ? 32 : 0;
      for (int r = 0; r < 32; ++r) {
        uint16_t raw;
        memcpy(&raw, blk + 4 + 3 * r, 2);
        if (raw == 0) continue;
        const int l = base + r;
        const double d = (double)raw * 0.002;
        const double xy = d * cv[l];
        x[n] = xy * sin_t[az > 36000 ? 36000 : az];
        y[n] = xy * cos_t[az > 36000 ? 36000 : az];
        z[n] = d * sv[l];
        distance[n] = d;
        intensity[n] = blk[4 + 3 * r + 2];
        laser_id[n] = (uint8_t)l;
        azimuth[n] = az;
        ms[n] = ts;
        ++n;
      }
    }
  }
  // frames 0..nf-1 are complete (what PacketDecoder::GetFrames returns); frame nf is still being filled
  if (nf + 1 < max_frames) frame_offsets[nf + 1] = n;
  *n_frames = nf;
  return n;
}
