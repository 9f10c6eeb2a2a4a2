// torch_ext.cpp — thin PyTorch glue over the C-ABI (include/igs_raster.h).
//
// Exposes the reference's pybind entry points with identical names, argument order and return
// tuples (DGR/ext.cpp:15-19, DGR/rasterize_points.h:18-81):
//   rasterize_gaussians, rasterize_gaussians_backward, mark_visible
// plus the batched V-view variants used by the multi-view refine step and a fused Adam step.
// Tensors are only used for memory ownership and stream lookup; all compute is in libigs_raster.so.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <string>
#include <tuple>
#include <vector>
#include "../../include/igs_raster.h"

namespace {

char* tensor_alloc(void* user, size_t bytes) {
  auto* t = reinterpret_cast<torch::Tensor*>(user);
  t->resize_({(long long)bytes});
  return reinterpret_cast<char*>(t->data_ptr());
}

const float* fptr(const torch::Tensor& t, const char* name) {
  if (!t.defined() || t.numel() == 0) return nullptr;
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (this rasterizer has no CPU path)");
  TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, " must be float32");
  TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
  return t.data_ptr<float>();
}

torch::Tensor contig(const torch::Tensor& t) {
  if (!t.defined() || t.numel() == 0) return t;
  return t.contiguous();
}

void check(int rc, const char* what) {
  if (rc != IGS_OK) {
    std::string msg = std::string(what) + " failed: " + igs_last_error();
    if (rc == IGS_ERR_INVALID || rc == IGS_ERR_UNSUPPORTED) throw std::runtime_error(msg);
    throw std::runtime_error(msg);
  }
}

using FwdRet = std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
                          torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>;

// Shared implementation for V >= 1.  `batched` only changes output shapes ([V,C,H,W] vs [C,H,W]).
FwdRet forward_impl(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                    const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                    float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                    const torch::Tensor& projmatrix, float tan_fovx, float tan_fovy, float kernel_size, int H, int W,
                    const torch::Tensor& sh, int degree, const torch::Tensor& campos, bool prefiltered,
                    bool require_coord, bool require_depth, bool debug, int V, bool batched,
                    const torch::Tensor& tile_rows = torch::Tensor(), int64_t instance_capacity = 0,
                    const torch::Tensor& status = torch::Tensor()) {
  if (means3D.ndimension() != 2 || means3D.size(1) != 3) {
    AT_ERROR("means3D must have dimensions (num_points, 3)");
  }
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (this rasterizer has no CPU path)");
  const c10::cuda::CUDAGuard guard(means3D.device());
  const int P = (int)means3D.size(0);
  auto fopt = means3D.options().dtype(torch::kFloat32);
  auto iopt = means3D.options().dtype(torch::kInt32);
  auto shape = [&](int C) { return batched ? std::vector<int64_t>{V, C, H, W} : std::vector<int64_t>{C, H, W}; };
  // One slab for the 15 output planes (the reference zero-fills 8 tensors separately).  With require_coord and
  // require_depth the blend kernel writes every pixel of every plane, so the slab needs no memset (1.6 GB at C3);
  // otherwise the planes that are not produced must read as zeros, like the reference's.
  const bool windowed = tile_rows.defined() && tile_rows.numel() != 0;
  const bool all_planes_written = require_coord && require_depth && P != 0;
  torch::Tensor slab = all_planes_written ? torch::empty({(int64_t)V * 15, H, W}, fopt)
                                          : torch::zeros({(int64_t)V * 15, H, W}, fopt);
  int64_t plane = (int64_t)H * W;
  auto view_of = [&](int64_t first, int C) {
    return slab.narrow(0, first, (int64_t)V * C).view(shape(C));
  };
  torch::Tensor out_color = view_of(0, 3);
  torch::Tensor out_coord = view_of((int64_t)V * 3, 3);
  torch::Tensor out_mcoord = view_of((int64_t)V * 6, 3);
  torch::Tensor out_normal = view_of((int64_t)V * 9, 3);
  torch::Tensor out_depth = view_of((int64_t)V * 12, 1);
  torch::Tensor out_mdepth = view_of((int64_t)V * 13, 1);
  torch::Tensor out_alpha = view_of((int64_t)V * 14, 1);
  (void)plane;
  torch::Tensor radii = batched ? torch::zeros({V, P}, iopt) : torch::zeros({P}, iopt);

  auto bopt = torch::TensorOptions().dtype(torch::kByte).device(means3D.device());
  torch::Tensor geomBuffer = torch::empty({0}, bopt);
  torch::Tensor binningBuffer = torch::empty({0}, bopt);
  torch::Tensor imgBuffer = torch::empty({0}, bopt);

  int rendered = 0;
  if (P != 0) {
    int M = 0;
    if (sh.defined() && sh.numel() != 0) M = (int)sh.size(1);
    torch::Tensor bg = contig(background), m3 = contig(means3D), col = contig(colors), op = contig(opacity),
                  sc = contig(scales), rot = contig(rotations), c3 = contig(cov3D_precomp), vm = contig(viewmatrix),
                  pm = contig(projmatrix), shc = contig(sh), cp = contig(campos);
    igs_forward_args a;
    memset(&a, 0, sizeof(a));
    a.P = P; a.D = degree; a.M = M; a.V = V; a.width = W; a.height = H;
    a.background = fptr(bg, "bg"); a.means3D = fptr(m3, "means3D"); a.shs = fptr(shc, "sh");
    a.colors_precomp = fptr(col, "colors_precomp"); a.opacities = fptr(op, "opacities");
    a.scales = fptr(sc, "scales"); a.scale_modifier = scale_modifier; a.rotations = fptr(rot, "rotations");
    a.cov3D_precomp = fptr(c3, "cov3D_precomp"); a.viewmatrix = fptr(vm, "viewmatrix");
    a.projmatrix = fptr(pm, "projmatrix"); a.cam_pos = fptr(cp, "campos");
    a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.kernel_size = kernel_size;
    a.prefiltered = prefiltered; a.require_coord = require_coord; a.require_depth = require_depth; a.debug = debug;
    a.out_color = out_color.data_ptr<float>(); a.out_coord = out_coord.data_ptr<float>();
    a.out_mcoord = out_mcoord.data_ptr<float>(); a.out_depth = out_depth.data_ptr<float>();
    a.out_mdepth = out_mdepth.data_ptr<float>(); a.out_alpha = out_alpha.data_ptr<float>();
    a.out_normal = out_normal.data_ptr<float>(); a.radii = radii.data_ptr<int>();
    torch::Tensor tr;
    if (windowed) {
      TORCH_CHECK(tile_rows.is_cuda() && tile_rows.scalar_type() == torch::kInt32 && tile_rows.numel() == (int64_t)V * 2,
                  "tile_rows must be a CUDA int32 tensor [V,2]");
      tr = tile_rows.contiguous();
      a.tile_rows = tr.data_ptr<int>();
    }
    TORCH_CHECK(instance_capacity >= 0 && instance_capacity <= 0x7FFFFFFF, "instance_capacity out of range");
    a.instance_capacity = (int)instance_capacity;
    if (status.defined() && status.numel() != 0) {
      TORCH_CHECK(status.is_cuda() && status.scalar_type() == torch::kInt32 && status.numel() >= 2 && status.is_contiguous(),
                  "status must be a contiguous CUDA int32 tensor with 2 elements");
      a.status = reinterpret_cast<unsigned int*>(status.data_ptr<int>());
    }
    TORCH_CHECK(a.background && a.opacities && a.viewmatrix && a.projmatrix && a.cam_pos, "missing required CUDA input");
    TORCH_CHECK(vm.numel() == (int64_t)V * 16 && pm.numel() == (int64_t)V * 16 && cp.numel() == (int64_t)V * 3,
                "camera tensors must hold ", V, " view(s)");
    if (a.colors_precomp == nullptr && a.shs == nullptr) throw std::runtime_error("For non-RGB, provide precomputed Gaussian colors!");
    cudaStream_t stream = at::cuda::getCurrentCUDAStream();
    check(igs_rasterize_forward(&a, tensor_alloc, &geomBuffer, tensor_alloc, &binningBuffer, tensor_alloc, &imgBuffer,
                                (void*)stream, &rendered),
          "rasterize_gaussians");
  }
  return std::make_tuple(rendered, out_color, out_coord, out_mcoord, out_alpha, out_normal, out_depth, out_mdepth, radii,
                         geomBuffer, binningBuffer, imgBuffer);
}

using BwdRet = std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
                          torch::Tensor, torch::Tensor>;

// Caller-provided destinations for the per-Gaussian gradients (contiguous float32 CUDA tensors of the right size).
struct BwdDest {
  torch::Tensor means3D, sh, opacity, scales, rotations;
};

BwdRet backward_impl(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                     const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                     float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                     const torch::Tensor& projmatrix, float tan_fovx, float tan_fovy, float kernel_size,
                     const torch::Tensor& dL_dout_color, const torch::Tensor& dL_dout_coord,
                     const torch::Tensor& dL_dout_mcoord, const torch::Tensor& dL_dout_depth,
                     const torch::Tensor& dL_dout_mdepth, const torch::Tensor& dL_dout_alpha,
                     const torch::Tensor& dL_dout_normal, const torch::Tensor& normalmap, const torch::Tensor& sh,
                     int degree, const torch::Tensor& campos, const torch::Tensor& geomBuffer, int R,
                     const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const torch::Tensor& alphas,
                     bool require_coord, bool require_depth, bool debug, int V, bool batched,
                     const BwdDest* dest = nullptr) {
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (this rasterizer has no CPU path)");
  const c10::cuda::CUDAGuard guard(means3D.device());
  const int P = (int)means3D.size(0);
  const int H = (int)alphas.size(-2), W = (int)alphas.size(-1);
  int M = 0;
  if (sh.defined() && sh.numel() != 0) M = (int)sh.size(1);
  auto fopt = means3D.options().dtype(torch::kFloat32);
  // `dest` lets the fused refine step have the per-Gaussian gradients written in place (packed gradient record)
  torch::Tensor dL_dmeans3D = dest ? dest->means3D : torch::empty({P, 3}, fopt);
  torch::Tensor dL_dmeans2D = batched ? torch::empty({V, P, 3}, fopt) : torch::empty({P, 3}, fopt);
  torch::Tensor dL_dcolors = torch::empty({P, 3}, fopt);
  torch::Tensor dL_dopacity = dest ? dest->opacity : torch::empty({P, 1}, fopt);
  torch::Tensor dL_dcov3D = torch::empty({P, 6}, fopt);
  torch::Tensor dL_dsh = dest ? dest->sh : torch::empty({P, M, 3}, fopt);
  torch::Tensor dL_dscales = dest ? dest->scales : torch::zeros({P, 3}, fopt);
  torch::Tensor dL_drotations = dest ? dest->rotations : torch::zeros({P, 4}, fopt);
  if (P != 0) {
    torch::Tensor bg = contig(background), m3 = contig(means3D), col = contig(colors), sc = contig(scales),
                  rot = contig(rotations), c3 = contig(cov3D_precomp), vm = contig(viewmatrix), pm = contig(projmatrix),
                  shc = contig(sh), cp = contig(campos), rd = contig(radii), al = contig(alphas), nm = contig(normalmap);
    torch::Tensor gc = contig(dL_dout_color), gco = contig(dL_dout_coord), gmc = contig(dL_dout_mcoord),
                  gd = contig(dL_dout_depth), gmd = contig(dL_dout_mdepth), ga = contig(dL_dout_alpha),
                  gn = contig(dL_dout_normal);
    if (!gc.defined() || gc.numel() == 0) gc = torch::zeros({(int64_t)V * 3, H, W}, fopt);
    igs_backward_args a;
    memset(&a, 0, sizeof(a));
    a.P = P; a.D = degree; a.M = M; a.V = V; a.R = R; a.width = W; a.height = H;
    a.background = fptr(bg, "bg"); a.means3D = fptr(m3, "means3D"); a.shs = fptr(shc, "sh");
    a.colors_precomp = fptr(col, "colors_precomp"); a.scales = fptr(sc, "scales"); a.scale_modifier = scale_modifier;
    a.rotations = fptr(rot, "rotations"); a.cov3D_precomp = fptr(c3, "cov3D_precomp");
    a.viewmatrix = fptr(vm, "viewmatrix"); a.projmatrix = fptr(pm, "projmatrix"); a.cam_pos = fptr(cp, "campos");
    a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.kernel_size = kernel_size;
    TORCH_CHECK(rd.is_cuda() && rd.scalar_type() == torch::kInt32, "radii must be a CUDA int32 tensor");
    a.radii = rd.data_ptr<int>(); a.alphas = fptr(al, "alphas"); a.normalmap = fptr(nm, "normalmap");
    a.geom_buffer = reinterpret_cast<char*>(geomBuffer.data_ptr());
    a.binning_buffer = reinterpret_cast<char*>(binningBuffer.data_ptr());
    a.image_buffer = reinterpret_cast<char*>(imageBuffer.data_ptr());
    a.dL_dout_color = fptr(gc, "grad_color"); a.dL_dout_coord = fptr(gco, "grad_coord");
    a.dL_dout_mcoord = fptr(gmc, "grad_mcoord"); a.dL_dout_depth = fptr(gd, "grad_depth");
    a.dL_dout_mdepth = fptr(gmd, "grad_mdepth"); a.dL_dout_alpha = fptr(ga, "grad_alpha");
    a.dL_dout_normal = fptr(gn, "grad_normal");
    a.dL_dmeans2D = dL_dmeans2D.data_ptr<float>(); a.dL_dcolors = dL_dcolors.data_ptr<float>();
    a.dL_dopacity = dL_dopacity.data_ptr<float>(); a.dL_dmeans3D = dL_dmeans3D.data_ptr<float>();
    a.dL_dcov3D = dL_dcov3D.data_ptr<float>(); a.dL_dsh = M ? dL_dsh.data_ptr<float>() : nullptr;
    a.dL_dscales = a.scales ? dL_dscales.data_ptr<float>() : nullptr;
    a.dL_drotations = a.scales ? dL_drotations.data_ptr<float>() : nullptr;
    a.require_coord = require_coord; a.require_depth = require_depth; a.debug = debug;
    auto bopt = torch::TensorOptions().dtype(torch::kByte).device(means3D.device());
    torch::Tensor temp = torch::empty({0}, bopt);
    cudaStream_t stream = at::cuda::getCurrentCUDAStream();
    check(igs_rasterize_backward(&a, tensor_alloc, &temp, (void*)stream), "rasterize_gaussians_backward");
  }
  return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
}

}  // namespace

// ---- reference-compatible entry points (single view) ----
FwdRet RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                              const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                              const float scale_modifier, const torch::Tensor& cov3D_precomp,
                              const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                              const float tan_fovy, const float kernel_size, const int image_height,
                              const int image_width, const torch::Tensor& sh, const int degree,
                              const torch::Tensor& campos, const bool prefiltered, const bool require_coord,
                              const bool require_depth, const bool debug) {
  return forward_impl(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                      projmatrix, tan_fovx, tan_fovy, kernel_size, image_height, image_width, sh, degree, campos,
                      prefiltered, require_coord, require_depth, debug, 1, false);
}

BwdRet RasterizeGaussiansBackwardCUDA(
    const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
    const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations, const float scale_modifier,
    const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
    const float tan_fovx, const float tan_fovy, const float kernel_size, const torch::Tensor& dL_dout_color,
    const torch::Tensor& dL_dout_coord, const torch::Tensor& dL_dout_mcoord, const torch::Tensor& dL_dout_depth,
    const torch::Tensor& dL_dout_mdepth, const torch::Tensor& dL_dout_alpha, const torch::Tensor& dL_dout_normal,
    const torch::Tensor& normalmap, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
    const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
    const torch::Tensor& alphas, const bool require_coord, const bool require_depth, const bool debug) {
  return backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                       projmatrix, tan_fovx, tan_fovy, kernel_size, dL_dout_color, dL_dout_coord, dL_dout_mcoord,
                       dL_dout_depth, dL_dout_mdepth, dL_dout_alpha, dL_dout_normal, normalmap, sh, degree, campos,
                       geomBuffer, R, binningBuffer, imageBuffer, alphas, require_coord, require_depth, debug, 1, false);
}

torch::Tensor markVisible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix) {
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (this rasterizer has no CPU path)");
  const c10::cuda::CUDAGuard guard(means3D.device());
  const int P = (int)means3D.size(0);
  torch::Tensor present = torch::full({P}, false, means3D.options().dtype(at::kBool));
  if (P != 0) {
    torch::Tensor m3 = means3D.contiguous(), vm = viewmatrix.contiguous(), pm = projmatrix.contiguous();
    check(igs_mark_visible(P, fptr(m3, "means3D"), fptr(vm, "viewmatrix"), fptr(pm, "projmatrix"),
                           reinterpret_cast<uint8_t*>(present.data_ptr<bool>()),
                           (void*)at::cuda::getCurrentCUDAStream().stream()),
          "mark_visible");
  }
  return present;
}

// ---- batched multi-view entry points (V cameras sharing H, W, tan_fov) ----
FwdRet RasterizeViewsCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                          const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                          const float scale_modifier, const torch::Tensor& cov3D_precomp,
                          const torch::Tensor& viewmatrices, const torch::Tensor& projmatrices, const float tan_fovx,
                          const float tan_fovy, const float kernel_size, const int image_height, const int image_width,
                          const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                          const bool prefiltered, const bool require_coord, const bool require_depth, const bool debug,
                          const c10::optional<torch::Tensor>& tile_rows_opt, int64_t instance_capacity,
                          const c10::optional<torch::Tensor>& status_opt) {
  const torch::Tensor tile_rows = tile_rows_opt.has_value() ? *tile_rows_opt : torch::Tensor();
  const torch::Tensor status = status_opt.has_value() ? *status_opt : torch::Tensor();
  TORCH_CHECK(viewmatrices.dim() == 3, "viewmatrices must be [V,4,4]");
  return forward_impl(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                      viewmatrices, projmatrices, tan_fovx, tan_fovy, kernel_size, image_height, image_width, sh, degree,
                      campos, prefiltered, require_coord, require_depth, debug, (int)viewmatrices.size(0), true,
                      tile_rows, instance_capacity, status);
}

BwdRet RasterizeViewsBackwardCUDA(
    const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
    const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations, const float scale_modifier,
    const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrices, const torch::Tensor& projmatrices,
    const float tan_fovx, const float tan_fovy, const float kernel_size, const torch::Tensor& dL_dout_color,
    const torch::Tensor& dL_dout_coord, const torch::Tensor& dL_dout_mcoord, const torch::Tensor& dL_dout_depth,
    const torch::Tensor& dL_dout_mdepth, const torch::Tensor& dL_dout_alpha, const torch::Tensor& dL_dout_normal,
    const torch::Tensor& normalmap, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
    const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
    const torch::Tensor& alphas, const bool require_coord, const bool require_depth, const bool debug) {
  return backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrices,
                       projmatrices, tan_fovx, tan_fovy, kernel_size, dL_dout_color, dL_dout_coord, dL_dout_mcoord,
                       dL_dout_depth, dL_dout_mdepth, dL_dout_alpha, dL_dout_normal, normalmap, sh, degree, campos,
                       geomBuffer, R, binningBuffer, imageBuffer, alphas, require_coord, require_depth, debug,
                       (int)viewmatrices.size(0), true);
}

// Decode the opaque buffers into reference-style arrays (parity tests).
std::vector<torch::Tensor> DecodeBuffers(int P, int V, int R, int H, int W, const torch::Tensor& geomBuffer,
                                         const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer) {
  const c10::cuda::CUDAGuard guard(geomBuffer.device());
  auto dev = geomBuffer.device();
  auto fo = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  auto io = torch::TensorOptions().dtype(torch::kInt32).device(dev);
  auto lo = torch::TensorOptions().dtype(torch::kInt64).device(dev);
  int tiles = ((W + 15) / 16) * ((H + 15) / 16);
  torch::Tensor keys = torch::zeros({R}, lo), plist = torch::zeros({R}, io), ranges = torch::zeros({V * tiles, 2}, io),
                ncontrib = torch::zeros({V, 2, H * W}, io), tiles_t = torch::zeros({V, P}, io),
                depths = torch::zeros({V, P}, fo), means2D = torch::zeros({V, P, 2}, fo),
                conic = torch::zeros({V, P, 4}, fo), rgb = torch::zeros({V, P, 3}, fo), cov3D = torch::zeros({P, 6}, fo),
                vp = torch::zeros({V, P, 3}, fo), cpl = torch::zeros({V, P, 6}, fo), rp = torch::zeros({V, P, 2}, fo),
                ts = torch::zeros({V, P}, fo), nrm = torch::zeros({V, P, 3}, fo);
  igs_decode_out o;
  o.keys = reinterpret_cast<uint64_t*>(keys.data_ptr<int64_t>());
  o.point_list = reinterpret_cast<uint32_t*>(plist.data_ptr<int>());
  o.ranges = reinterpret_cast<uint32_t*>(ranges.data_ptr<int>());
  o.n_contrib = reinterpret_cast<uint32_t*>(ncontrib.data_ptr<int>());
  o.tiles_touched = reinterpret_cast<uint32_t*>(tiles_t.data_ptr<int>());
  o.depths = depths.data_ptr<float>(); o.means2D = means2D.data_ptr<float>(); o.conic_opacity = conic.data_ptr<float>();
  o.rgb = rgb.data_ptr<float>(); o.cov3D = cov3D.data_ptr<float>(); o.view_points = vp.data_ptr<float>();
  o.camera_planes = cpl.data_ptr<float>(); o.ray_planes = rp.data_ptr<float>(); o.ts = ts.data_ptr<float>();
  o.normals = nrm.data_ptr<float>();
  check(igs_decode_binning(P, V, R, W, H, reinterpret_cast<const char*>(geomBuffer.data_ptr()),
                           R > 0 ? reinterpret_cast<const char*>(binningBuffer.data_ptr()) : nullptr,
                           reinterpret_cast<const char*>(imageBuffer.data_ptr()), &o,
                           (void*)at::cuda::getCurrentCUDAStream().stream()),
        "decode_buffers");
  return {keys, plist, ranges, ncontrib, tiles_t, depths, means2D, conic, rgb, cov3D, vp, cpl, rp, ts, nrm};
}

// Work counters of the blend traversal for bench.py's roofline: int64 [4] = {block-level hits, E' evaluated pixel pairs,
// C contributing pixel pairs, instances staged} (igs_count_blend_pairs).
torch::Tensor CountBlendPairs(int P, int V, int R, int H, int W, const torch::Tensor& geomBuffer,
                              const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer) {
  const c10::cuda::CUDAGuard guard(geomBuffer.device());
  torch::Tensor counts = torch::zeros({4}, torch::TensorOptions().dtype(torch::kInt64).device(geomBuffer.device()));
  check(igs_count_blend_pairs(P, V, R, W, H, reinterpret_cast<const char*>(geomBuffer.data_ptr()),
                              R > 0 ? reinterpret_cast<const char*>(binningBuffer.data_ptr()) : nullptr,
                              reinterpret_cast<const char*>(imageBuffer.data_ptr()),
                              reinterpret_cast<unsigned long long*>(counts.data_ptr<int64_t>()),
                              (void*)at::cuda::getCurrentCUDAStream().stream()),
        "count_blend_pairs");
  return counts;
}

void AdamStep(torch::Tensor& param, const torch::Tensor& grad, torch::Tensor& exp_avg, torch::Tensor& exp_avg_sq,
              double lr, double beta1, double beta2, double eps, int64_t step) {
  TORCH_CHECK(param.is_cuda() && grad.is_cuda() && exp_avg.is_cuda() && exp_avg_sq.is_cuda(), "Adam tensors must be CUDA");
  TORCH_CHECK(param.is_contiguous() && grad.is_contiguous() && exp_avg.is_contiguous() && exp_avg_sq.is_contiguous(),
              "Adam tensors must be contiguous");
  TORCH_CHECK(param.scalar_type() == torch::kFloat32 && grad.scalar_type() == torch::kFloat32, "Adam tensors must be float32");
  TORCH_CHECK(param.numel() == grad.numel() && param.numel() == exp_avg.numel() && param.numel() == exp_avg_sq.numel(),
              "Adam tensor sizes differ");
  const c10::cuda::CUDAGuard guard(param.device());
  check(igs_adam_step(param.data_ptr<float>(), grad.data_ptr<float>(), exp_avg.data_ptr<float>(),
                      exp_avg_sq.data_ptr<float>(), (size_t)param.numel(), (float)lr, (float)beta1, (float)beta2,
                      (float)eps, (int)step, (void*)at::cuda::getCurrentCUDAStream().stream()),
        "adam_step");
}

std::tuple<torch::Tensor, torch::Tensor> RefineLoss(const torch::Tensor& pred, const torch::Tensor& gt, double lambda_l1,
                                                    double grad_scale, const std::vector<int64_t>& loss_rows) {
  TORCH_CHECK(pred.is_cuda() && gt.is_cuda(), "refine_loss needs CUDA tensors (there is no CPU path)");
  TORCH_CHECK(pred.dim() == 4 && pred.size(1) == 3 && pred.sizes() == gt.sizes(), "pred/gt must be [V,3,H,W]");
  TORCH_CHECK(pred.scalar_type() == torch::kFloat32 && gt.scalar_type() == torch::kFloat32, "float32 only");
  const c10::cuda::CUDAGuard guard(pred.device());
  torch::Tensor p = pred.contiguous(), g = gt.contiguous();
  torch::Tensor loss = torch::zeros({}, p.options());
  torch::Tensor dpred = torch::empty_like(p);
  torch::Tensor temp = torch::empty({0}, torch::TensorOptions().dtype(torch::kByte).device(p.device()));
  TORCH_CHECK(loss_rows.empty() || (int64_t)loss_rows.size() == 2 * p.size(0), "loss_rows must hold 2 ints per view");
  std::vector<int> rows(loss_rows.begin(), loss_rows.end());
  check(igs_refine_loss(p.data_ptr<float>(), g.data_ptr<float>(), (int)p.size(0), (int)p.size(2), (int)p.size(3),
                        (float)lambda_l1, (float)grad_scale, rows.empty() ? nullptr : rows.data(), loss.data_ptr<float>(),
                        dpred.data_ptr<float>(), tensor_alloc, &temp,
                        (void*)at::cuda::getCurrentCUDAStream().stream()),
        "refine_loss");
  return std::make_tuple(loss, dpred);
}

// (loss_sum, dL_dmap0, dL_dmap1, dL_dnormal) of the depth-normal consistency loss (igs_depth_normal_loss); maps are
// [V,1,H,W] / [V,H,W] depths or [V,3,H,W] coords, normal [V,3,H,W]
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor> DepthNormalLoss(
    const torch::Tensor& map0, const torch::Tensor& map1, const torch::Tensor& normal, bool use_coord, double tan_fovx,
    double tan_fovy, double depth_ratio, double grad_scale) {
  TORCH_CHECK(map0.is_cuda() && map1.is_cuda() && normal.is_cuda(), "depth_normal_loss needs CUDA tensors (there is no CPU path)");
  TORCH_CHECK(normal.dim() == 4 && normal.size(1) == 3, "normal must be [V,3,H,W]");
  TORCH_CHECK(map0.scalar_type() == torch::kFloat32 && map1.scalar_type() == torch::kFloat32 &&
              normal.scalar_type() == torch::kFloat32, "float32 only");
  const int64_t V = normal.size(0), H = normal.size(2), W = normal.size(3);
  const int64_t want = V * (use_coord ? 3 : 1) * H * W;
  TORCH_CHECK(map0.numel() == want && map1.numel() == want, "maps must hold V*", use_coord ? 3 : 1, "*H*W floats");
  const c10::cuda::CUDAGuard guard(normal.device());
  torch::Tensor a = map0.contiguous(), b = map1.contiguous(), n = normal.contiguous();
  torch::Tensor loss = torch::zeros({}, n.options());
  torch::Tensor da = torch::empty_like(a), db = torch::empty_like(b), dn = torch::empty_like(n);
  torch::Tensor temp = torch::empty({0}, torch::TensorOptions().dtype(torch::kByte).device(n.device()));
  check(igs_depth_normal_loss(use_coord ? 1 : 0, a.data_ptr<float>(), b.data_ptr<float>(), n.data_ptr<float>(), (int)V, (int)H,
                              (int)W, (float)tan_fovx, (float)tan_fovy, (float)depth_ratio, (float)grad_scale,
                              loss.data_ptr<float>(), da.data_ptr<float>(), db.data_ptr<float>(), dn.data_ptr<float>(),
                              tensor_alloc, &temp, (void*)at::cuda::getCurrentCUDAStream().stream()),
        "depth_normal_loss");
  return std::make_tuple(loss, da, db, dn);
}

// Adam over the packed parameter record: one launch for all groups.
void AdamStepGroups(torch::Tensor& param, const torch::Tensor& grad, torch::Tensor& exp_avg, torch::Tensor& exp_avg_sq,
                    const std::vector<int64_t>& sizes, const std::vector<double>& lrs, double beta1, double beta2,
                    double eps, int64_t step, const c10::optional<torch::Tensor>& skip_opt,
                    const c10::optional<torch::Tensor>& step_dev_opt) {
  TORCH_CHECK(param.is_cuda() && grad.is_cuda() && exp_avg.is_cuda() && exp_avg_sq.is_cuda(), "Adam tensors must be CUDA");
  TORCH_CHECK(param.is_contiguous() && grad.is_contiguous() && exp_avg.is_contiguous() && exp_avg_sq.is_contiguous(),
              "Adam tensors must be contiguous");
  TORCH_CHECK(param.scalar_type() == torch::kFloat32 && grad.scalar_type() == torch::kFloat32, "Adam tensors must be float32");
  TORCH_CHECK(sizes.size() == lrs.size() && sizes.size() <= 8, "at most 8 groups, one learning rate each");
  int64_t total = 0;
  std::vector<size_t> sz;
  std::vector<float> lr;
  for (size_t k = 0; k < sizes.size(); k++) { total += sizes[k]; sz.push_back((size_t)sizes[k]); lr.push_back((float)lrs[k]); }
  TORCH_CHECK(total == param.numel() && total == grad.numel() && total == exp_avg.numel() && total == exp_avg_sq.numel(),
              "group sizes must add up to the buffer size");
  const c10::cuda::CUDAGuard guard(param.device());
  if (skip_opt.has_value() || step_dev_opt.has_value()) {
    const unsigned int* skip = nullptr;
    int* step_dev = nullptr;
    if (skip_opt.has_value()) {
      TORCH_CHECK(skip_opt->is_cuda() && skip_opt->scalar_type() == torch::kInt32 && skip_opt->numel() >= 1, "skip must be a CUDA int32 tensor");
      skip = reinterpret_cast<const unsigned int*>(skip_opt->data_ptr<int>());
    }
    if (step_dev_opt.has_value()) {
      TORCH_CHECK(step_dev_opt->is_cuda() && step_dev_opt->scalar_type() == torch::kInt32 && step_dev_opt->numel() == 1, "step_dev must be a CUDA int32 scalar");
      step_dev = step_dev_opt->data_ptr<int>();
    }
    check(igs_adam_step_groups_guarded(param.data_ptr<float>(), grad.data_ptr<float>(), exp_avg.data_ptr<float>(),
                                       exp_avg_sq.data_ptr<float>(), (int)sz.size(), sz.data(), lr.data(), (float)beta1,
                                       (float)beta2, (float)eps, (int)step, skip, step_dev,
                                       (void*)at::cuda::getCurrentCUDAStream().stream()),
          "adam_step_groups_guarded");
    return;
  }
  check(igs_adam_step_groups(param.data_ptr<float>(), grad.data_ptr<float>(), exp_avg.data_ptr<float>(),
                             exp_avg_sq.data_ptr<float>(), (int)sz.size(), sz.data(), lr.data(), (float)beta1,
                             (float)beta2, (float)eps, (int)step, (void*)at::cuda::getCurrentCUDAStream().stream()),
        "adam_step_groups");
}

// (scales [P,3], rotations [P,4], opacities [P,1]) = activations of the raw refine parameters
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> ActivateForward(const torch::Tensor& log_scales,
                                                                        const torch::Tensor& raw_rot,
                                                                        const torch::Tensor& logit) {
  TORCH_CHECK(log_scales.is_cuda(), "activations need CUDA tensors (there is no CPU path)");
  const c10::cuda::CUDAGuard guard(log_scales.device());
  const int P = (int)log_scales.size(0);
  torch::Tensor ls = log_scales.contiguous(), rr = raw_rot.contiguous(), lg = logit.contiguous();
  torch::Tensor sc = torch::empty({P, 3}, ls.options()), rot = torch::empty({P, 4}, ls.options()),
                op = torch::empty({P, 1}, ls.options());
  check(igs_activate_forward(P, fptr(ls, "log_scales"), fptr(rr, "raw_rot"), fptr(lg, "logit"), sc.data_ptr<float>(),
                             rot.data_ptr<float>(), op.data_ptr<float>(),
                             (void*)at::cuda::getCurrentCUDAStream().stream()),
        "activate_forward");
  return std::make_tuple(sc, rot, op);
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> ActivateBackward(
    const torch::Tensor& scales, const torch::Tensor& raw_rot, const torch::Tensor& opac, const torch::Tensor& g_scales,
    const torch::Tensor& g_rot, const torch::Tensor& g_opac) {
  TORCH_CHECK(scales.is_cuda(), "activations need CUDA tensors (there is no CPU path)");
  const c10::cuda::CUDAGuard guard(scales.device());
  const int P = (int)scales.size(0);
  torch::Tensor sc = scales.contiguous(), rr = raw_rot.contiguous(), op = opac.contiguous(), gs = g_scales.contiguous(),
                gr = g_rot.contiguous(), go = g_opac.contiguous();
  torch::Tensor d_ls = torch::empty({P, 3}, sc.options()), d_rr = torch::empty({P, 4}, sc.options()),
                d_lg = torch::empty({P, 1}, sc.options());
  check(igs_activate_backward(P, fptr(sc, "scales"), fptr(rr, "raw_rot"), fptr(op, "opac"), fptr(gs, "g_scales"),
                              fptr(gr, "g_rot"), fptr(go, "g_opac"), d_ls.data_ptr<float>(), d_rr.data_ptr<float>(),
                              d_lg.data_ptr<float>(), (void*)at::cuda::getCurrentCUDAStream().stream()),
        "activate_backward");
  return std::make_tuple(d_ls, d_rr, d_lg);
}

// This rank's share of one refine iteration with NO autograd graph and no per-op Python:
//   activations -> batched forward -> fused loss (dL/dcolor scaled by 1/total_views) -> backward ->
//   activation chain rule, with every per-Gaussian gradient written in place into the packed record `grad`.
// `flat` / `grad` are the [P*59] parameter / gradient records; offs = element offsets of
// (xyz, rotation, shs, opacity, scaling) inside them.  Returns (loss_sum/total_views [device scalar],
// num_rendered, radii [V,P], dL_dmeans2D [V,P,3]).
// Mirrors infer_batch.py:279-324 for V views at once (see igs_b200/refine.py).
std::tuple<torch::Tensor, int64_t, torch::Tensor, torch::Tensor, torch::Tensor> RefineGrad(
    const torch::Tensor& flat, torch::Tensor& grad, const std::vector<int64_t>& offs, int64_t P64,
    const torch::Tensor& background, const torch::Tensor& viewmatrices, const torch::Tensor& projmatrices,
    const torch::Tensor& campos, double tan_fovx, double tan_fovy, double kernel_size, double scale_modifier, int H,
    int W, int degree, const torch::Tensor& gt, int64_t total_views, double lambda_l1, bool require_coord,
    bool require_depth, bool debug, const c10::optional<torch::Tensor>& tile_rows_opt,
    const std::vector<int64_t>& loss_rows, int64_t instance_capacity, const c10::optional<torch::Tensor>& status_opt) {
  const torch::Tensor tile_rows = tile_rows_opt.has_value() ? *tile_rows_opt : torch::Tensor();
  TORCH_CHECK(flat.is_cuda() && grad.is_cuda(), "refine_grad needs CUDA tensors (there is no CPU path)");
  // {true instance count, overflow flag} of this call's forward, on the device (capacity mode: see igs_forward_args)
  torch::Tensor status = status_opt.has_value() ? *status_opt
                                                : torch::zeros({2}, torch::TensorOptions().dtype(torch::kInt32).device(flat.device()));
  TORCH_CHECK(flat.is_contiguous() && grad.is_contiguous() && flat.numel() == grad.numel(), "flat/grad must be contiguous and equal-sized");
  TORCH_CHECK(flat.scalar_type() == torch::kFloat32 && grad.scalar_type() == torch::kFloat32, "float32 only");
  TORCH_CHECK(offs.size() == 5, "offs = (xyz, rotation, shs, opacity, scaling)");
  TORCH_CHECK(viewmatrices.dim() == 3, "viewmatrices must be [V,4,4]");
  const c10::cuda::CUDAGuard guard(flat.device());
  const int P = (int)P64;
  const int V = (int)viewmatrices.size(0);
  TORCH_CHECK(gt.dim() == 4 && gt.size(0) == V && gt.size(1) == 3 && gt.size(2) == H && gt.size(3) == W, "gt must be [V,3,H,W]");
  TORCH_CHECK(total_views >= 1, "total_views must be >= 1");
  auto slice = [&](const torch::Tensor& buf, int k, int64_t per, std::vector<int64_t> shape) {
    return buf.narrow(0, offs[k], (int64_t)P * per).view(shape);
  };
  torch::Tensor xyz = slice(flat, 0, 3, {P, 3}), raw_rot = slice(flat, 1, 4, {P, 4}), shs = slice(flat, 2, 48, {P, 16, 3}),
                logit = slice(flat, 3, 1, {P, 1}), log_sc = slice(flat, 4, 3, {P, 3});
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  // activated parameters and their gradients: one 16P-float slab (rotations first: float4-aligned)
  torch::Tensor slab = torch::empty({(int64_t)P * 16}, flat.options());
  torch::Tensor rot = slab.narrow(0, 0, (int64_t)P * 4).view({P, 4}), g_rot = slab.narrow(0, (int64_t)P * 4, (int64_t)P * 4).view({P, 4}),
                sc = slab.narrow(0, (int64_t)P * 8, (int64_t)P * 3).view({P, 3}), g_sc = slab.narrow(0, (int64_t)P * 11, (int64_t)P * 3).view({P, 3}),
                op = slab.narrow(0, (int64_t)P * 14, P).view({P, 1}), g_op = slab.narrow(0, (int64_t)P * 15, P).view({P, 1});
  check(igs_activate_forward(P, log_sc.data_ptr<float>(), raw_rot.data_ptr<float>(), logit.data_ptr<float>(),
                             sc.data_ptr<float>(), rot.data_ptr<float>(), op.data_ptr<float>(), (void*)stream),
        "activate_forward");
  torch::Tensor none;
  FwdRet f = forward_impl(background, xyz, none, op, sc, rot, (float)scale_modifier, none, viewmatrices, projmatrices,
                          (float)tan_fovx, (float)tan_fovy, (float)kernel_size, H, W, shs, degree, campos, false,
                          require_coord, require_depth, debug, V, true, tile_rows, instance_capacity, status);
  const int rendered = std::get<0>(f);
  const torch::Tensor& color = std::get<1>(f);
  torch::Tensor loss = torch::zeros({}, flat.options());
  torch::Tensor dcolor = torch::empty_like(color);
  torch::Tensor temp = torch::empty({0}, torch::TensorOptions().dtype(torch::kByte).device(flat.device()));
  torch::Tensor gtc = gt.contiguous();
  TORCH_CHECK(gtc.is_cuda() && gtc.scalar_type() == torch::kFloat32, "gt must be a float32 CUDA tensor");
  const float gscale = 1.0f / (float)total_views;
  TORCH_CHECK(loss_rows.empty() || (int64_t)loss_rows.size() == 2 * (int64_t)V, "loss_rows must hold 2 ints per view");
  std::vector<int> rows(loss_rows.begin(), loss_rows.end());
  check(igs_refine_loss(color.data_ptr<float>(), gtc.data_ptr<float>(), V, H, W, (float)lambda_l1, gscale,
                        rows.empty() ? nullptr : rows.data(), loss.data_ptr<float>(), dcolor.data_ptr<float>(), tensor_alloc, &temp, (void*)stream),
        "refine_loss");
  BwdDest dest;
  dest.means3D = slice(grad, 0, 3, {P, 3});
  dest.sh = slice(grad, 2, 48, {P, 16, 3});
  dest.opacity = g_op; dest.scales = g_sc; dest.rotations = g_rot;
  BwdRet b = backward_impl(background, xyz, std::get<8>(f), none, sc, rot, (float)scale_modifier, none, viewmatrices,
                           projmatrices, (float)tan_fovx, (float)tan_fovy, (float)kernel_size, dcolor, none, none, none,
                           none, none, none, std::get<5>(f), shs, degree, campos, std::get<9>(f), rendered,
                           std::get<10>(f), std::get<11>(f), std::get<4>(f), require_coord, require_depth, debug, V, true,
                           &dest);
  check(igs_activate_backward(P, sc.data_ptr<float>(), raw_rot.data_ptr<float>(), op.data_ptr<float>(),
                              g_sc.data_ptr<float>(), g_rot.data_ptr<float>(), g_op.data_ptr<float>(),
                              slice(grad, 4, 3, {P, 3}).data_ptr<float>(), slice(grad, 1, 4, {P, 4}).data_ptr<float>(),
                              slice(grad, 3, 1, {P, 1}).data_ptr<float>(), (void*)stream),
        "activate_backward");
  loss.mul_(gscale);
  return std::make_tuple(loss, (int64_t)rendered, std::get<8>(f), std::get<0>(b), status);
}

// Fused exchange (igs_fused_reduce_adam): grad_mc / param_mc are MULTICAST addresses (integers from the symmetric-memory
// handle), everything else ordinary tensors of this rank.  Returns nothing; loss_out [1] float and skipped_out [1] int32
// are written on the device.
void FusedReduceAdam(int64_t grad_mc, int64_t param_mc, const torch::Tensor& param, torch::Tensor& exp_avg,
                     torch::Tensor& exp_avg_sq, int64_t n, int64_t begin, int64_t end, const std::vector<int64_t>& sizes,
                     const std::vector<double>& lrs, double beta1, double beta2, double eps, int64_t step,
                     torch::Tensor& loss_out, torch::Tensor& skipped_out) {
  TORCH_CHECK(param.is_cuda() && exp_avg.is_cuda() && exp_avg_sq.is_cuda() && loss_out.is_cuda() && skipped_out.is_cuda(),
              "fused_reduce_adam needs CUDA tensors (there is no CPU path)");
  TORCH_CHECK(grad_mc != 0 && param_mc != 0, "multicast addresses are null (no NVLS multicast support on this system?)");
  TORCH_CHECK(param.numel() >= n && exp_avg.numel() >= n && exp_avg_sq.numel() >= n, "buffers shorter than n");
  TORCH_CHECK(loss_out.scalar_type() == torch::kFloat32 && skipped_out.scalar_type() == torch::kInt32, "loss_out float32, skipped_out int32");
  TORCH_CHECK(sizes.size() == lrs.size() && !sizes.empty() && sizes.size() <= 8, "1..8 groups");
  std::vector<size_t> sz(sizes.begin(), sizes.end());
  std::vector<float> lr(lrs.begin(), lrs.end());
  const c10::cuda::CUDAGuard guard(param.device());
  check(igs_fused_reduce_adam(reinterpret_cast<const float*>(grad_mc), reinterpret_cast<float*>(param_mc), param.data_ptr<float>(),
                              exp_avg.data_ptr<float>(), exp_avg_sq.data_ptr<float>(), (size_t)n, (size_t)begin, (size_t)end,
                              (int)sz.size(), sz.data(), lr.data(), (float)beta1, (float)beta2, (float)eps, (int)step,
                              loss_out.data_ptr<float>(), reinterpret_cast<unsigned int*>(skipped_out.data_ptr<int>()),
                              (void*)at::cuda::getCurrentCUDAStream().stream()),
        "fused_reduce_adam");
}

void DensifyStats(const torch::Tensor& g2d, const torch::Tensor& radii, torch::Tensor& grad_accum, torch::Tensor& denom,
                  torch::Tensor& max_radii, double grad_scale) {
  TORCH_CHECK(g2d.is_cuda() && radii.is_cuda() && grad_accum.is_cuda(), "densify_stats needs CUDA tensors (there is no CPU path)");
  TORCH_CHECK(g2d.dim() == 3 && g2d.size(2) == 3 && radii.dim() == 2, "dL_dmeans2D must be [V,P,3] and radii [V,P]");
  TORCH_CHECK(radii.scalar_type() == torch::kInt32, "radii must be int32");
  const int V = (int)g2d.size(0), P = (int)g2d.size(1);
  TORCH_CHECK(radii.size(0) == V && radii.size(1) == P && grad_accum.numel() == P && denom.numel() == P && max_radii.numel() == P,
              "size mismatch");
  TORCH_CHECK(grad_accum.is_contiguous() && denom.is_contiguous() && max_radii.is_contiguous(), "statistics must be contiguous");
  const c10::cuda::CUDAGuard guard(g2d.device());
  torch::Tensor g = g2d.contiguous(), r = radii.contiguous();
  check(igs_densify_stats(P, V, fptr(g, "dL_dmeans2D"), r.data_ptr<int>(), (float)grad_scale, grad_accum.data_ptr<float>(),
                          denom.data_ptr<float>(), max_radii.data_ptr<float>(),
                          (void*)at::cuda::getCurrentCUDAStream().stream()),
        "densify_stats");
}

// (param_new, exp_avg_new, exp_avg_sq_new) = packed records of P_new Gaussians gathered from the old ones
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> DensifyGather(
    const torch::Tensor& param, const torch::Tensor& exp_avg, const torch::Tensor& exp_avg_sq, int64_t P_old,
    const torch::Tensor& src, const torch::Tensor& kind, const torch::Tensor& samples, double split_div,
    const std::vector<int64_t>& off_old, const std::vector<int64_t>& off_new) {
  TORCH_CHECK(param.is_cuda() && exp_avg.is_cuda() && exp_avg_sq.is_cuda() && src.is_cuda() && kind.is_cuda(),
              "densify_gather needs CUDA tensors (there is no CPU path)");
  TORCH_CHECK(src.scalar_type() == torch::kInt32 && kind.scalar_type() == torch::kUInt8, "src int32, kind uint8");
  TORCH_CHECK(off_old.size() == 5 && off_new.size() == 5, "offsets of (xyz, rotation, shs, opacity, scaling)");
  TORCH_CHECK(param.numel() == P_old * 59 && exp_avg.numel() == P_old * 59 && exp_avg_sq.numel() == P_old * 59,
              "old records must hold P_old * 59 floats");
  const int64_t P_new = src.numel();
  TORCH_CHECK(kind.numel() == P_new, "kind / src size mismatch");
  const c10::cuda::CUDAGuard guard(param.device());
  torch::Tensor pn = torch::empty({P_new * 59}, param.options()), mn = torch::empty({P_new * 59}, param.options()),
                vn = torch::empty({P_new * 59}, param.options());
  torch::Tensor sc = src.contiguous(), kc = kind.contiguous(), sm = samples.defined() ? samples.contiguous() : samples;
  if (sm.defined() && sm.numel() != 0) TORCH_CHECK(sm.numel() == P_new * 3 && sm.is_cuda(), "samples must be [P_new,3] on the device");
  size_t oo[5], on[5];
  for (int k = 0; k < 5; k++) { oo[k] = (size_t)off_old[k]; on[k] = (size_t)off_new[k]; }
  torch::Tensor po = param.contiguous(), mo = exp_avg.contiguous(), vo = exp_avg_sq.contiguous();
  check(igs_densify_gather((int)P_old, (int)P_new, sc.data_ptr<int>(), kc.data_ptr<uint8_t>(), fptr(sm, "samples"),
                           (float)split_div, oo, on, po.data_ptr<float>(), mo.data_ptr<float>(), vo.data_ptr<float>(),
                           pn.data_ptr<float>(), mn.data_ptr<float>(), vn.data_ptr<float>(),
                           (void*)at::cuda::getCurrentCUDAStream().stream()),
        "densify_gather");
  return std::make_tuple(pn, mn, vn);
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("refine_loss", &RefineLoss, pybind11::arg("pred"), pybind11::arg("gt"), pybind11::arg("lambda_l1"),
        pybind11::arg("grad_scale") = 1.0, pybind11::arg("loss_rows") = std::vector<int64_t>());
  m.def("refine_grad", &RefineGrad, pybind11::arg("flat"), pybind11::arg("grad"), pybind11::arg("offs"), pybind11::arg("P"),
        pybind11::arg("background"), pybind11::arg("viewmatrices"), pybind11::arg("projmatrices"), pybind11::arg("campos"),
        pybind11::arg("tan_fovx"), pybind11::arg("tan_fovy"), pybind11::arg("kernel_size"), pybind11::arg("scale_modifier"),
        pybind11::arg("H"), pybind11::arg("W"), pybind11::arg("degree"), pybind11::arg("gt"), pybind11::arg("total_views"),
        pybind11::arg("lambda_l1"), pybind11::arg("require_coord"), pybind11::arg("require_depth"), pybind11::arg("debug"),
        pybind11::arg("tile_rows") = pybind11::none(), pybind11::arg("loss_rows") = std::vector<int64_t>(),
        pybind11::arg("instance_capacity") = 0, pybind11::arg("status") = pybind11::none());
  m.def("depth_normal_loss", &DepthNormalLoss, pybind11::arg("map0"), pybind11::arg("map1"), pybind11::arg("normal"),
        pybind11::arg("use_coord"), pybind11::arg("tan_fovx"), pybind11::arg("tan_fovy"), pybind11::arg("depth_ratio") = 0.6,
        pybind11::arg("grad_scale") = 1.0);
  m.def("activate_forward", &ActivateForward);
  m.def("activate_backward", &ActivateBackward);
  m.def("adam_step_groups", &AdamStepGroups, pybind11::arg("param"), pybind11::arg("grad"), pybind11::arg("exp_avg"),
        pybind11::arg("exp_avg_sq"), pybind11::arg("sizes"), pybind11::arg("lrs"), pybind11::arg("beta1"), pybind11::arg("beta2"),
        pybind11::arg("eps"), pybind11::arg("step"), pybind11::arg("skip") = pybind11::none(),
        pybind11::arg("step_dev") = pybind11::none());
  m.def("count_blend_pairs", &CountBlendPairs);
  m.def("fused_reduce_adam", &FusedReduceAdam);
  m.def("densify_stats", &DensifyStats, pybind11::arg("dL_dmeans2D"), pybind11::arg("radii"), pybind11::arg("grad_accum"),
        pybind11::arg("denom"), pybind11::arg("max_radii2D"), pybind11::arg("grad_scale") = 1.0);
  m.def("densify_gather", &DensifyGather);
  m.def("rasterize_gaussians", &RasterizeGaussiansCUDA);
  m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA);
  m.def("mark_visible", &markVisible);
  m.def("rasterize_views", &RasterizeViewsCUDA, pybind11::arg("background"), pybind11::arg("means3D"), pybind11::arg("colors"),
        pybind11::arg("opacity"), pybind11::arg("scales"), pybind11::arg("rotations"), pybind11::arg("scale_modifier"),
        pybind11::arg("cov3D_precomp"), pybind11::arg("viewmatrices"), pybind11::arg("projmatrices"),
        pybind11::arg("tan_fovx"), pybind11::arg("tan_fovy"), pybind11::arg("kernel_size"), pybind11::arg("image_height"),
        pybind11::arg("image_width"), pybind11::arg("sh"), pybind11::arg("degree"), pybind11::arg("campos"),
        pybind11::arg("prefiltered"), pybind11::arg("require_coord"), pybind11::arg("require_depth"),
        pybind11::arg("debug"), pybind11::arg("tile_rows") = pybind11::none(), pybind11::arg("instance_capacity") = 0,
        pybind11::arg("status") = pybind11::none());
  m.def("rasterize_views_backward", &RasterizeViewsBackwardCUDA);
  m.def("decode_buffers", &DecodeBuffers);
  m.def("adam_step", &AdamStep);
  m.def("abi_version", []() { return igs_abi_version(); });
  m.def("stage_timing_enable", [](bool on) { igs_stage_timing_enable(on ? 1 : 0); });
  m.def("stage_timing_read", [](bool reset) {
    double ms[64]; unsigned long long cnt[64];
    int n = igs_stage_timing_read(ms, cnt, 64, reset ? 1 : 0);
    std::vector<std::tuple<std::string, double, int64_t>> out;
    for (int i = 0; i < n; i++) out.emplace_back(std::string(igs_stage_name(i)), ms[i], (int64_t)cnt[i]);
    return out;
  });
  m.def("kernel_launches", []() { return (int64_t)igs_kernel_launches(); });
}
