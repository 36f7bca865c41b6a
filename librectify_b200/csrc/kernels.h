// kernels.h — host-callable launchers of the sm_100a kernels (internal; the public C ABI is include/librectify_b200.h)
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "common.cuh"

namespace lrb {

// k1_grad.cu
void launch_grad(const CUtensorMap & tmap_img, const Planes & P, FrameCtrl * ctrl, int n_slots, int tma_x, int tma_y, cudaStream_t st);
int  grad_tile_w();
int  grad_tile_h();
// k2_nms.cu
void launch_frame_init(const Seeds & S, FrameCtrl * ctrl, int n_slots, cudaStream_t st);
void launch_nms(const CUtensorMap & tmap_mag, const CUtensorMap & tmap_bin, const Planes & P, const Seeds & S,
                FrameCtrl * ctrl, int n_slots, int tma_x, int tma_y, cudaStream_t st);
void launch_seed_order(const Planes & P, const Seeds & S, FrameCtrl * ctrl, int n_slots, int seeds_hint, cudaStream_t st);
void nms_tile_dims(int * mw, int * mh, int * bw, int * bh);
// k3_replay.cu: canonical order of exact-magnitude ties = what libstdc++ std::sort leaves (replay of introsort's partitions)
void launch_seed_replay(const Planes & P, const Seeds & S, FrameCtrl * ctrl, int n_slots, int depth_override, bool fixup, cudaStream_t st);
// k4_flood.cu
int launch_flood_rounds(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots, int seeds_hint,
                         int round_begin, int round_end, unsigned epoch_first, cudaStream_t st);
void launch_flood_serial(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots,
                         unsigned epoch_serial, cudaStream_t st);
// k5_fit.cu
void launch_fit(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots, int rounds_total, cudaStream_t st);
void launch_labels(const Planes & P, const Comps & K, FrameCtrl * ctrl, int * labels, int slot, cudaStream_t st);
// k6_ransac.cu
void launch_filter_model(const Comps & K, const Lines & L, FrameCtrl * ctrl, int n_slots, float min_length, int rounds_total, bool refined, cudaStream_t st);
// ev_score: nullptr, or 2 * kMaxModels events recorded before / after the k_score launch of each solve (per-kernel roofline)
void launch_ransac(const Lines & L, FrameCtrl * ctrl, int n_slots, const unsigned * stream, unsigned long long stream_len,
                   int T, cudaStream_t st, cudaEvent_t * ev_score = nullptr);
// k_refine.cu
int  launch_refine(const Comps & K, FrameCtrl * ctrl, int n_slots, cudaStream_t st);      // returns the kernels launched
const float * refine_output(const Comps & K, int slot);                                    // merged raw segments of a slot
// k7_sampler.cu
void launch_sampler(const Lines & L, FrameCtrl * ctrl, int n_slots, const unsigned * stream, unsigned long long stream_len,
                    int solve, int T, cudaStream_t st);
void launch_sampler_only(const Lines & L, FrameCtrl * ctrl, const unsigned * stream, unsigned long long stream_len, int T, cudaStream_t st);
void launch_sampler_parallel_only(const Lines & L, FrameCtrl * ctrl, const unsigned * stream, unsigned long long stream_len, int T, cudaStream_t st);
// mt_stream.cu
void launch_mt_stream(unsigned seed, unsigned * stream, unsigned long long n, cudaStream_t st);
// ingest.cu
void launch_ingest(const float * src, long long src_pitch, long long src_frame_stride, float * dst, int dst_pitch,
                   size_t dst_stride, int w, int h, int n_slots, cudaStream_t st);
void launch_ingest_u8(const unsigned char * src, long long src_pitch, long long src_frame_stride, int channels, float * dst,
                      int dst_pitch, size_t dst_stride, int w, int h, int n_slots, cudaStream_t st);
void launch_fill_u32(unsigned * p, size_t n, unsigned v, cudaStream_t st);
void launch_mt_raw(unsigned seed, unsigned * out, unsigned long long n, cudaStream_t st);
// ubench.cu: FP32-pipe peaks (mode 0: FFMA, 1: alternating FMUL / FADD); out: blocks * 256 floats
void launch_ubench_fp32(int mode, float * out, int blocks, int iters, cudaStream_t st);
// k_hough.cu
size_t hough_out_bytes();
int hough_draws();
int hough_hypotheses();
void launch_hough_direct(const float * seg7, int n_lines, const int * idx, int m, const unsigned * raw, int n_raw,
                         float * h, float * anchor, float * dir, float * len, int * draws, int * hyp_cell, float * hyp_val,
                         float * weights, int * inl, void * out, cudaStream_t st);
// k_aux.cu: fit_vanishing_point / assign_to_group on the device
void launch_aux_model(const float * seg7, int n, unsigned * box4, float * h, float * len, float * norm3, cudaStream_t st);
void launch_aux_common_points(const float * seg7, int n, const float * h, const float * len, const float * norm3,
                              const int * groups, int n_groups, bool all, float * out, cudaStream_t st);
void launch_aux_assign(float * fresh7, int n_fresh, const int * groups, const float * vps, int n_groups, float thr, cudaStream_t st);
// allocation-free fit_vanishing_point on device-resident lines: scratch in a static __device__ array of k_aux.cu (<= kAuxNoAllocMaxLines
// lines); enqueues on `st` and returns; 0, or a negative code with a message in err
constexpr int kAuxNoAllocMaxLines = 65536;
int aux_fit_vanishing_point_noalloc(int device, const float * d_seg7, int n, int group, float * d_out3, cudaStream_t st,
                                    char * err, size_t err_len);
// k_prosac.cu: PROSAC_Estimator::solve (reference src/prosac.h:104-299) on a model + weights left on the device by launch_hough_direct
size_t prosac_out_bytes();
void launch_prosac(const float * h, const float * anchor, const float * dir, const float * len, int n_lines, const int * idx, int m,
                   const float * weights, const unsigned * stream, unsigned long long stream_len, float tol,
                   int * sidx, unsigned char * flags, int * inl, void * out, cudaStream_t st);
// k_warp.cu
struct WarpMatrix { double m[9]; };    // destination -> source map, row-major (what cv::warpPerspective uses after inverting H)
void launch_warp_perspective_u8(const unsigned char * src, int sw, int sh, long long spitch, int channels, unsigned char * dst, int dw, int dh,
                                long long dpitch, const WarpMatrix & M, cudaStream_t st);
void launch_resize_area_f32(const float * src, long long spitch, float * dst, int dw, int dh, long long dpitch,
                            const int * xofs, const int * xsi, const float * xalpha, const int * yofs, const int * ysi, const float * yalpha,
                            cudaStream_t st);
void launch_resize_area_fast_f32(const float * src, int sw, int sh, long long spitch, float * dst, int dw, int dh, long long dpitch, int scale,
                                 cudaStream_t st);

} // namespace lrb
