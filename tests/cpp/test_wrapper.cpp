// Drives include/lmpc.hpp (lmpc::ConvexMpcEngine, lmpc::ShardedEngine) through liblmpc.so.  Prints one
// line per check; exit code 0 = all good, 3 = no usable GPU (the engine refused to construct: there is
// no CPU fallback).  include/lmpc_legged_adapter.hpp is exercised by oracle/adapter_ref_main.cpp, which is
// compiled against the reference's real headers (oracle/Makefile), not here.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <vector>

#include <array>

#include "lmpc.hpp"

// argv: shard <in.bin> <out.bin> <preset> <n_shards>: solve the batch of in.bin (int32 B, int32 H, then x0,
// rot, foot_pos_abs, cmd as doubles and contact as bytes) through lmpc::ShardedEngine; shards sit
// on distinct GPUs when the box has them, otherwise all on device 0.  Writes grf [B][12] + status [B].
static int run_shard(char** argv) {
  std::ifstream f(argv[2], std::ios::binary);
  std::int32_t B = 0, H = 0;
  f.read(reinterpret_cast<char*>(&B), 4);
  f.read(reinterpret_cast<char*>(&H), 4);
  std::vector<double> x0(12 * B), rot(9 * B), foot(12 * B), cmd(9 * B), grf(12 * B);
  std::vector<std::uint8_t> contact((size_t)4 * H * B);
  std::vector<std::int32_t> status(B, -1);
  f.read(reinterpret_cast<char*>(x0.data()), 8 * x0.size());
  f.read(reinterpret_cast<char*>(rot.data()), 8 * rot.size());
  f.read(reinterpret_cast<char*>(foot.data()), 8 * foot.size());
  f.read(reinterpret_cast<char*>(cmd.data()), 8 * cmd.size());
  f.read(reinterpret_cast<char*>(contact.data()), contact.size());
  if (!f) { std::printf("short input file\n"); return 2; }
  const int G = std::atoi(argv[5]);
  try {
    lmpc::ConvexMpcEngine probe(lmpc::preset(argv[4]), 1, 0);  // throws without a GPU
    std::vector<int> devs;
    for (int g = 0; g < G; ++g) {
      int d = g;
      try { lmpc::ConvexMpcEngine e(lmpc::preset(argv[4]), 1, d); } catch (const lmpc::Error&) { d = 0; }
      devs.push_back(d);
    }
    lmpc::ShardedEngine eng(lmpc::preset(argv[4]), B, devs);
    lmpc_batch_in in{};
    lmpc_batch_out out{};
    in.x0 = x0.data(); in.rot = rot.data(); in.foot_pos_abs = foot.data(); in.cmd = cmd.data(); in.contact = contact.data();
    out.grf = grf.data(); out.status = status.data();
    // a batch beyond the engine's capacity must come back as an exception from the argument check,
    // before any shard thread starts (the old header threw with joinable threads alive)
    int overflow_rejected = 0;
    try { eng.solve(B + 3 * G, in, out); } catch (const lmpc::Error& e) { overflow_rejected = e.code() == LMPC_ERR_INVALID_ARG; }
    const long long bad = eng.solve(B, in, out);
    std::ofstream o(argv[3], std::ios::binary);
    o.write(reinterpret_cast<const char*>(grf.data()), 8 * grf.size());
    o.write(reinterpret_cast<const char*>(status.data()), 4 * status.size());
    std::printf("shards %d devices", eng.shards());
    for (int d : devs) std::printf(" %d", d);
    std::printf(" unconverged %lld overflow_rejected %d\n", bad, overflow_rejected);
    return 0;
  } catch (const lmpc::Error& e) {
    std::printf("lmpc::Error %d: %s\n", (int)e.code(), e.what());
    return e.code() == LMPC_ERR_NO_DEVICE ? 3 : 2;
  }
}

int main(int argc, char** argv) {
  if (argc >= 6 && !std::strcmp(argv[1], "shard")) return run_shard(argv);
  // One reference-preset tick of a standing A1 in trot phase 0 through lmpc::ConvexMpcEngine, fed the way
  // the adapter feeds it (x0, rot, foot_pos_abs, cmd, contact[H][4]); prints the 12 GRFs.
  const double q[12] = {60.0, 100.0, 0.0, 0.0, 0.0, 450.0, 0.15, 0.15, 100.0, 3.0, 3.0, 5.0};
  double r[12];
  for (int i = 0; i < 12; ++i) r[i] = 1e-4;
  const double fh[4][3] = {{0.17, 0.17, -0.3}, {0.17, -0.17, -0.3}, {-0.17, 0.17, -0.3}, {-0.17, -0.17, -0.3}};
  try {
    lmpc::ConvexMpcEngine ref(q, r, "reference_a1", 1);
    ref.set_warm_start(true, 1);
    lmpc::RobotTick t;
    t.x0[5] = 0.3; t.cmd[2] = 0.3;
    t.rot[0] = t.rot[4] = t.rot[8] = 1.0;
    for (int l = 0; l < 4; ++l)
      for (int a = 0; a < 3; ++a) t.foot_pos_abs[3 * l + a] = fh[l][a];
    const double ph0[4] = {0.0, 0.0, 0.0, 0.0};
    const auto c0 = ref.contact_table(LMPC_GAIT_TROT, ph0, 3.5);
    std::memcpy(t.contact.data(), c0.data(), c0.size());
    std::array<double, 12> grf{};
    for (int tick = 0; tick < 3; ++tick) grf = ref.compute_grfs(t);  // ticks 2 and 3 start from the stored faces
    std::printf("grf");
    for (int i = 0; i < 12; ++i) std::printf(" %.17g", grf[i]);
    std::printf("\n");
    // the gait helper and a stand on the north-star preset
    lmpc::ConvexMpcEngine eng(q, r, "northstar_a1", 4);
    const double ph[4] = {0.1, 0.1, 0.1, 0.1};
    const auto c = eng.contact_table(LMPC_GAIT_STAND, ph, 3.5);
    std::memcpy(t.contact.data(), c.data(), c.size());
    int st = -1;
    const auto g = eng.compute_grfs(t, &st);
    std::printf("stand status %d fz %.17g %.17g %.17g %.17g\n", st, g[2], g[5], g[8], g[11]);
    return (st == 0 && std::fabs(g[2] - g[11]) < 1e-6) ? 0 : 1;
  } catch (const lmpc::Error& e) {
    std::printf("lmpc::Error %d: %s\n", (int)e.code(), e.what());
    return e.code() == LMPC_ERR_NO_DEVICE ? 3 : 2;
  }
}
