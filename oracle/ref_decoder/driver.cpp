// oracle/ref_decoder/driver.cpp — TEST INFRASTRUCTURE: runs the REFERENCE's own prebuilt Velodyne packet
// decoder (lib/libPacketDecoder.so, declared in include/PacketDecoder.h; it is what src/pcapReader.cpp:79-102
// drives) on a file of raw 1206-byte packets and dumps the frames it produces.  Used only to pin the decode
// oracle and to write the golden fixtures under tests/golden/ (tools in this directory); nothing on the
// product path, and nothing at GPU-box run time, touches it.  Built by `make -C oracle ref` into oracle/_ref/.
//
// usage: ref_decoder <16db.xml> <packets.bin> <frames.bin> [max_frames]
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include <fstream>
#include <iostream>
#include <iterator>

#include "PacketDecoder.h"

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  std::ifstream in(argv[2], std::ios::binary);
  std::string all((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
  const size_t n_pkt = all.size() / 1206;
  PacketDecoder dec;
  dec.SetMaxNumberOfFrames(argc > 4 ? (unsigned)atoi(argv[4]) : 100000u);
  dec.SetCorrectionsFile(argv[1]);
  for (size_t p = 0; p < n_pkt; ++p) {
    std::string pkt = all.substr(p * 1206, 1206);
    unsigned int len = 1206;
    dec.DecodePacket(&pkt, &len);
  }
  std::deque<PacketDecoder::HDLFrame> frames = dec.GetFrames();
  FILE* out = fopen(argv[3], "wb");
  if (!out) return 3;
  const uint64_t nf = frames.size();
  fwrite(&nf, 8, 1, out);
  for (const auto& f : frames) {
    const uint64_t n = f.x.size();
    fwrite(&n, 8, 1, out);
    fwrite(f.x.data(), 8, n, out);
    fwrite(f.y.data(), 8, n, out);
    fwrite(f.z.data(), 8, n, out);
    fwrite(f.distance.data(), 8, n, out);
    fwrite(f.intensity.data(), 1, n, out);
    fwrite(f.laser_id.data(), 1, n, out);
    fwrite(f.azimuth.data(), 2, n, out);
    fwrite(f.ms_from_top_of_hour.data(), 4, n, out);
  }
  fclose(out);
  fprintf(stderr, "packets %zu frames %llu\n", n_pkt, (unsigned long long)nf);
  return 0;
}
