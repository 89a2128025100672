// Internal description of the two MLP layer graphs behind nb_model_t.
#pragma once
#include <string>
#include <vector>

#include "common.cuh"

namespace nb {

// One dense layer: weight [out, in] row-major (nn.Linear convention) at w_off, bias [out] at b_off,
// both offsets in floats into the caller's flat parameter buffer.
struct Lin {
  std::string name;
  int in = 0, out = 0;
  int64_t w_off = 0, b_off = 0;
};

// Both graphs are instances of one shape:
//   trunk : depth layers of `width`; layer 0 reads the xyz encoding (Cx); the layer after `skip`
//           reads concat([xyz encoding, h]) (TF order, nerf/run_nerf_helpers.py:97-100);
//           ReLU after every trunk layer except layer 0 when !act0 (FlexibleNeRFModel.layer1,
//           nerf-pytorch/nerf/models.py:238).
//   heads : alpha = Lin(h -> 1); feat = Lin(h -> feat_width) [+ReLU when feat_act];
//           dir = ReLU(Lin([feat, dir encoding] -> width/2)); rgb = Lin(dir -> 3);  raw = [rgb, alpha].
//           Without view directions: raw = Lin(h -> 4).
struct Graph {
  int kind = 0, depth = 0, width = 0, n_freq_xyz = 0, n_freq_dir = 0, use_viewdirs = 0, skip = -1;
  int Cx = 0, Cd = 0;
  bool act0 = true, feat_act = false;
  int feat_width = 0, dir_width = 0;
  std::vector<Lin> lins;
  // indices into lins
  std::vector<int> trunk;
  int alpha = -1, feat = -1, dir = -1, rgb = -1, out4 = -1;
  int64_t n_params = 0;
};

struct PackedBf16;  // mlp_tc.cu

}  // namespace nb

struct nb_model {
  nb::Graph g;
  const float* params = nullptr;  // borrowed flat fp32 parameters (device)
  int64_t n_params_set = 0;
  nb::PackedBf16* packed = nullptr;  // owned bf16 tile images for the tcgen05 path
};

namespace nb {
void packed_free(PackedBf16* p);
int packed_build(nb_model* m, cudaStream_t stream);  // (re)build bf16 images from m->params
}  // namespace nb
