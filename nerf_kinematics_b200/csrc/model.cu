// nb_model_t: the two MLP layer graphs of the path described as one
// parametrised shape, plus the flat-parameter layout the C-ABI exposes (include/nerf_b200.h).
//
// The reference builds these with Keras (nerf/run_nerf_helpers.py:80-118, init_nerf_model) and
// torch.nn (nerf-pytorch/nerf/models.py:186-231, FlexibleNeRFModel.__init__).  Nothing here is a
// kernel; it is the host-side graph description the fp32 and tcgen05 MLP kernels are driven by.
#include <string>

#include "model.h"

namespace nb {

static void add_lin(Graph& g, const std::string& name, int in, int out, int* slot) {
  Lin l;
  l.name = name;
  l.in = in;
  l.out = out;
  l.w_off = g.n_params;
  l.b_off = l.w_off + (int64_t)in * out;
  g.n_params = l.b_off + out;
  *slot = (int)g.lins.size();
  g.lins.push_back(l);
}

static int build_graph(Graph& g) {
  g.Cx = 3 + 6 * g.n_freq_xyz;
  g.Cd = g.use_viewdirs ? 3 + 6 * g.n_freq_dir : 0;
  const bool tf = g.kind == NB_GRAPH_NERF;
  g.act0 = tf;               // FlexibleNeRFModel.layer1 has no activation (models.py:238)
  g.feat_act = !tf;          // Flexible: relu(fc_feat) (models.py:248); TF bottleneck is linear (helpers:104)
  g.feat_width = tf ? 256 : g.width;  // Dense(256) is hard-wired in init_nerf_model (helpers:104)
  g.dir_width = g.width / 2;
  if (!tf) g.skip = -1;      // the Flexible skip never fires at its shipped depth and crashes otherwise
  int fin = g.Cx;
  for (int i = 0; i < g.depth; ++i) {
    int slot;
    std::string nm = tf ? ("pts_linears." + std::to_string(i))
                        : (i == 0 ? std::string("layer1") : "layers_xyz." + std::to_string(i - 1));
    add_lin(g, nm, fin, g.width, &slot);
    g.trunk.push_back(slot);
    fin = (i == g.skip) ? g.width + g.Cx : g.width;
  }
  if (g.use_viewdirs) {
    if (tf) {
      add_lin(g, "alpha_linear", fin, 1, &g.alpha);
      add_lin(g, "bottleneck_linear", fin, g.feat_width, &g.feat);
      add_lin(g, "views_linears.0", g.feat_width + g.Cd, g.dir_width, &g.dir);
      add_lin(g, "rgb_linear", g.dir_width, 3, &g.rgb);
    } else {
      add_lin(g, "fc_feat", fin, g.feat_width, &g.feat);
      add_lin(g, "fc_alpha", fin, 1, &g.alpha);
      add_lin(g, "layers_dir.0", g.feat_width + g.Cd, g.dir_width, &g.dir);
      add_lin(g, "fc_rgb", g.dir_width, 3, &g.rgb);
    }
  } else {
    add_lin(g, tf ? "output_linear" : "fc_out", fin, 4, &g.out4);
  }
  return NB_OK;
}

int64_t graph_flops_per_point(const Graph& g) {
  int64_t mac = 0;
  for (const Lin& l : g.lins) mac += (int64_t)l.in * l.out;
  return 2 * mac;
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_model_create(int graph, int depth, int width, int n_freq_xyz, int n_freq_dir, int use_viewdirs, int skip_layer,
                    nb_model_t** out) {
  NB_REQUIRE(out != nullptr, "nb_model_create: out is null");
  *out = nullptr;
  NB_REQUIRE(graph == NB_GRAPH_NERF || graph == NB_GRAPH_FLEXIBLE, "nb_model_create: unknown graph %d", graph);
  NB_REQUIRE(depth >= 1 && depth <= 32 && width >= 8 && width <= 1024 && (width % 2) == 0,
             "nb_model_create: depth %d / width %d out of range", depth, width);
  NB_REQUIRE(n_freq_xyz >= 0 && n_freq_xyz <= 16 && n_freq_dir >= 0 && n_freq_dir <= 16,
             "nb_model_create: frequency counts out of range");
  NB_REQUIRE(skip_layer < depth - 1 || skip_layer < 0 || graph == NB_GRAPH_NERF,
             "nb_model_create: skip_layer %d out of range", skip_layer);
  nb_model* m = new nb_model();
  m->g.kind = graph;
  m->g.depth = depth;
  m->g.width = width;
  m->g.n_freq_xyz = n_freq_xyz;
  m->g.n_freq_dir = n_freq_dir;
  m->g.use_viewdirs = use_viewdirs ? 1 : 0;
  m->g.skip = skip_layer >= 0 ? skip_layer : -1;
  build_graph(m->g);
  *out = m;
  return NB_OK;
}

int nb_model_destroy(nb_model_t* m) {
  if (m == nullptr) return NB_OK;
  if (m->packed != nullptr) packed_free(m->packed);
  delete m;
  return NB_OK;
}

int64_t nb_model_param_count(const nb_model_t* m) { return m ? m->g.n_params : 0; }
int nb_model_num_tensors(const nb_model_t* m) { return m ? 2 * (int)m->g.lins.size() : 0; }

int nb_model_tensor_info(const nb_model_t* m, int index, char* name, int name_cap, int64_t* rows, int64_t* cols,
                         int64_t* offset_floats) {
  NB_REQUIRE(m != nullptr, "nb_model_tensor_info: null model");
  NB_REQUIRE(index >= 0 && index < 2 * (int)m->g.lins.size(), "nb_model_tensor_info: index %d out of range", index);
  const Lin& l = m->g.lins[index / 2];
  const bool bias = (index & 1) != 0;
  if (name != nullptr && name_cap > 0) snprintf(name, (size_t)name_cap, "%s.%s", l.name.c_str(), bias ? "bias" : "weight");
  if (rows) *rows = l.out;
  if (cols) *cols = bias ? 0 : l.in;
  if (offset_floats) *offset_floats = bias ? l.b_off : l.w_off;
  return NB_OK;
}

int nb_model_set_params(nb_model_t* m, const float* flat, int64_t n) {
  NB_REQUIRE(m != nullptr && flat != nullptr, "nb_model_set_params: null argument");
  NB_REQUIRE(n == m->g.n_params, "nb_model_set_params: got %lld floats, the graph has %lld parameters", (long long)n,
             (long long)m->g.n_params);
  m->params = flat;
  m->n_params_set = n;
  return NB_OK;
}

int nb_model_pack(nb_model_t* m, int precision, nb_stream_t stream) {
  NB_REQUIRE(m != nullptr, "nb_model_pack: null model");
  NB_REQUIRE(m->params != nullptr, "nb_model_pack: nb_model_set_params has not been called");
  if (precision == NB_PREC_FP32) return NB_OK;  // the fp32 kernels read the flat buffer directly
  NB_REQUIRE(precision == NB_PREC_BF16, "nb_model_pack: unknown precision %d", precision);
  return packed_build(m, as_stream(stream));
}

int64_t nb_model_flops_per_point(const nb_model_t* m) { return m ? graph_flops_per_point(m->g) : 0; }

}  // extern "C"
