// Test-only harness: runs the K5 frame arithmetic (csrc/fk_math.cuh, the same source the CUDA kernel inlines) on
// the host so that tests/test_fk_host.py can check it against the float64 oracle without a GPU.
// usage: fk_host_harness in.bin out.bin ; in = int32 N, int32 H, float dt, qpos[N*H*35], ref[H*35]
//        out = reward[N*H], terms[N*H*4] (pose, vel, ee, com squared-error sums), body_pos[N*H*45]
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../deepmimic_diffusion_mujoco_b200/csrc/fk_math.cuh"

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 3;
  int N, H; float dt;
  if (fread(&N, 4, 1, f) != 1 || fread(&H, 4, 1, f) != 1 || fread(&dt, 4, 1, f) != 1) return 4;
  std::vector<float> q((size_t)N * H * 35), ref((size_t)H * 35);
  if (fread(q.data(), 4, q.size(), f) != q.size() || fread(ref.data(), 4, ref.size(), f) != ref.size()) return 4;
  fclose(f);
  using namespace dm::fk;
  std::vector<float> feat((size_t)H * kFeatStride, 0.f);
  const float inv_dt = 1.f / dt;
  for (int h = 0; h < H; ++h)
    frame_pass<true, false>(&ref[(size_t)h * 35], &ref[(size_t)(h > 0 ? h - 1 : 0) * 35], inv_dt, nullptr, &feat[(size_t)h * kFeatStride], nullptr, nullptr);
  std::vector<float> reward((size_t)N * H), terms((size_t)N * H * 4), bp((size_t)N * H * 45);
  for (int n = 0; n < N; ++n)
    for (int h = 0; h < H; ++h) {
      const size_t fi = (size_t)n * H + h;
      const float* qq = &q[fi * 35];
      const float* qp = h > 0 ? qq - 35 : qq;
      float acc[4];
      frame_pass<false, true>(qq, qp, inv_dt, &feat[(size_t)h * kFeatStride], nullptr, &bp[fi * 45], acc);
      for (int k = 0; k < 4; ++k) terms[fi * 4 + k] = acc[k];
      reward[fi] = 0.65f * fast_exp(-2.f * acc[0]) + 0.1f * fast_exp(-0.1f * acc[1]) + 0.15f * fast_exp(-40.f * acc[2]) +
                   0.1f * fast_exp(-10.f * acc[3]);
    }
  f = fopen(argv[2], "wb");
  fwrite(reward.data(), 4, reward.size(), f);
  fwrite(terms.data(), 4, terms.size(), f);
  fwrite(bp.data(), 4, bp.size(), f);
  fclose(f);
  return 0;
}
