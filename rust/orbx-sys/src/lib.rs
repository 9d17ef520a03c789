//! Raw FFI to liborbx: GENERATED from include/orbx.h by scripts/gen_rust_sys.py - do not edit.
//! Every declaration of the header appears here once; the safe wrappers live in the `orbx` crate.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct orbx_keypoint {
    pub x: f32,
    pub y: f32,
    pub size: f32,
    pub angle: f32,
    pub response: f32,
    pub octave: i32,
    pub class_id: i32,
}
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct orbx_dmatch {
    pub query_idx: i32,
    pub train_idx: i32,
    pub img_idx: i32,
    pub distance: f32,
}
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct orbx_params {
    pub nfeatures: i32,
    pub scale_factor: f32,
    pub nlevels: i32,
    pub edge_threshold: i32,
    pub first_level: i32,
    pub wta_k: i32,
    pub score_type: i32,
    pub patch_size: i32,
    pub fast_threshold: i32,
}
pub enum orbx_detector {}
pub enum orbx_matcher {}
pub enum orbx_map {}
pub enum orbx_pose_estimator {}
pub enum orbx_triangulator {}
pub enum orbx_sequence {}
pub enum orbx_comm {}

pub const ORBX_OK: c_int = 0;
pub const ORBX_E_BADARG: c_int = -1;
pub const ORBX_E_CAPACITY: c_int = -2;
pub const ORBX_E_CUDA: c_int = -3;
pub const ORBX_E_NOMEM: c_int = -4;
pub const ORBX_E_DATA: c_int = -5;
pub const ORBX_HARRIS_SCORE: c_int = 0;
pub const ORBX_STAGE_PYRAMID: c_int = 0;
pub const ORBX_STAGE_FAST: c_int = 1;
pub const ORBX_STAGE_HARRIS: c_int = 2;
pub const ORBX_STAGE_SELECT: c_int = 3;
pub const ORBX_STAGE_ORIENT: c_int = 4;
pub const ORBX_STAGE_BLUR: c_int = 5;
pub const ORBX_STAGE_BRIEF: c_int = 6;
pub const ORBX_STAGE_COUNT: c_int = 7;
pub const ORBX_POSE_OK: c_int = 0;
pub const ORBX_POSE_INSUFFICIENT_POINTS: c_int = 1;
pub const ORBX_POSE_NO_MODEL: c_int = 2;
pub const ORBX_POSE_TOO_FEW_INLIERS: c_int = 3;

extern "C" {
    pub fn orbx_default_params(p: *mut orbx_params);
    pub fn orbx_detector_create(params: *const orbx_params, device: c_int, max_w: c_int, max_h: c_int, max_batch: c_int, out: *mut *mut orbx_detector) -> c_int;
    pub fn orbx_detector_destroy(d: *mut orbx_detector);
    pub fn orbx_detector_max_features(d: *const orbx_detector) -> c_int;
    pub fn orbx_detect(d: *mut orbx_detector, img: *const u8, w: c_int, h: c_int, pitch: usize, kps: *mut orbx_keypoint, cap: c_int, n: *mut c_int) -> c_int;
    pub fn orbx_detect_and_compute(d: *mut orbx_detector, img: *const u8, w: c_int, h: c_int, pitch: usize, kps: *mut orbx_keypoint, desc: *mut u8, cap: c_int, n: *mut c_int) -> c_int;
    pub fn orbx_detect_and_compute_batch(d: *mut orbx_detector, imgs: *const u8, w: c_int, h: c_int, pitch: usize, frame_stride: usize, batch: c_int, kps: *mut orbx_keypoint, desc: *mut u8, cap_per_frame: c_int, n_per_frame: *mut c_int) -> c_int;
    pub fn orbx_detect_and_compute_batch_bgr(d: *mut orbx_detector, imgs: *const u8, w: c_int, h: c_int, pitch: usize, frame_stride: usize, batch: c_int, kps: *mut orbx_keypoint, desc: *mut u8, cap_per_frame: c_int, n_per_frame: *mut c_int) -> c_int;
    pub fn orbx_detector_sync(d: *mut orbx_detector) -> c_int;
    pub fn orbx_detector_set_stream(d: *mut orbx_detector, stream: *mut c_void) -> c_int;
    pub fn orbx_detector_set_graph(d: *mut orbx_detector, enable: c_int) -> c_int;
    pub fn orbx_detector_level_info(d: *const orbx_detector, level: c_int, blurred: c_int, w: *mut c_int, h: *mut c_int, pitch: *mut usize, frame_stride: *mut usize, dev_ptr: *mut *const u8) -> c_int;
    pub fn orbx_detector_read_level(d: *mut orbx_detector, level: c_int, blurred: c_int, frame: c_int, host_out: *mut u8, out_pitch: usize) -> c_int;
    pub fn orbx_detector_candidates(d: *mut orbx_detector, frame: c_int, level: c_int, out: *mut u32, cap: c_int, n: *mut c_int) -> c_int;
    pub fn orbx_detector_set_timing(d: *mut orbx_detector, enable: c_int) -> c_int;
    pub fn orbx_detector_timing(d: *mut orbx_detector, ms_per_stage: *mut f32, launches_per_stage: *mut c_int, reset: c_int) -> c_int;
    pub fn orbx_matcher_create(device: c_int, cross_check: c_int, out: *mut *mut orbx_matcher) -> c_int;
    pub fn orbx_matcher_destroy(m: *mut orbx_matcher);
    pub fn orbx_matcher_sync(m: *mut orbx_matcher) -> c_int;
    pub fn orbx_matcher_set_stream(m: *mut orbx_matcher, stream: *mut c_void) -> c_int;
    pub fn orbx_match(m: *mut orbx_matcher, q: *const u8, nq: c_int, t: *const u8, nt: c_int, out: *mut orbx_dmatch, n: *mut c_int) -> c_int;
    pub fn orbx_knn_match2(m: *mut orbx_matcher, q: *const u8, nq: c_int, t: *const u8, nt: c_int, out: *mut orbx_dmatch, n_per_q: *mut c_int) -> c_int;
    pub fn orbx_knn_match2_batch(m: *mut orbx_matcher, q: *const u8, q_stride: usize, nq_dev: *const c_int, t: *const u8, t_stride: usize, nt_dev: *const c_int, npairs: c_int, max_rows: c_int, out: *mut orbx_dmatch, n_per_q: *mut c_int) -> c_int;
    pub fn orbx_ratio_test(m: *mut orbx_matcher, knn2: *const orbx_dmatch, n_per_q: *const c_int, nq: c_int, ratio: f32, out: *mut orbx_dmatch, n: *mut c_int) -> c_int;
    pub fn orbx_ratio_test_batch(m: *mut orbx_matcher, knn2: *const orbx_dmatch, n_per_q: *const c_int, nq_dev: *const c_int, npairs: c_int, max_rows: c_int, ratio: f32, out: *mut orbx_dmatch, n_out_dev: *mut c_int) -> c_int;
    pub fn orbx_filter_good_matches(m: *mut orbx_matcher, in_: *const orbx_dmatch, n: c_int, ratio: f32, out: *mut orbx_dmatch, n_out: *mut c_int) -> c_int;
    pub fn orbx_filter_good_matches_batch(m: *mut orbx_matcher, in_: *const orbx_dmatch, n_dev: *const c_int, npairs: c_int, max_rows: c_int, ratio: f32, out: *mut orbx_dmatch, n_out_dev: *mut c_int) -> c_int;
    pub fn orbx_match_batch(m: *mut orbx_matcher, q: *const u8, q_stride: usize, nq_dev: *const c_int, t: *const u8, t_stride: usize, nt_dev: *const c_int, npairs: c_int, max_rows: c_int, out: *mut orbx_dmatch, n_out_dev: *mut c_int) -> c_int;
    pub fn orbx_knn_match2_static(device: c_int, q: *const u8, nq: c_int, t: *const u8, nt: c_int, out: *mut orbx_dmatch, n_per_q: *mut c_int, stream: *mut c_void) -> c_int;
    pub fn orbx_extract_matched_points(m: *mut orbx_matcher, kp1: *const orbx_keypoint, n1: c_int, kp2: *const orbx_keypoint, n2: c_int, matches: *const orbx_dmatch, n: c_int, pts1: *mut f32, pts2: *mut f32) -> c_int;
    pub fn orbx_extract_matched_points_batch(m: *mut orbx_matcher, kp1: *const orbx_keypoint, n1_dev: *const c_int, kp2: *const orbx_keypoint, n2_dev: *const c_int, kp_stride: usize, matches: *const orbx_dmatch, n_dev: *const c_int, npairs: c_int, max_rows: c_int, pts1: *mut f32, pts2: *mut f32, bad_dev: *mut c_int) -> c_int;
    pub fn orbx_matcher_set_mode(m: *mut orbx_matcher, mode: c_int) -> c_int;
    pub fn orbx_matcher_set_timing(m: *mut orbx_matcher, enable: c_int) -> c_int;
    pub fn orbx_matcher_timing(m: *mut orbx_matcher, ms: *mut f32, launches: *mut c_int, reset: c_int) -> c_int;
    pub fn orbx_map_create(device: c_int, capacity: c_int, fx: f64, fy: f64, cx: f64, cy: f64, out: *mut *mut orbx_map) -> c_int;
    pub fn orbx_map_destroy(m: *mut orbx_map);
    pub fn orbx_map_size(m: *const orbx_map) -> c_int;
    pub fn orbx_map_clear(m: *mut orbx_map) -> c_int;
    pub fn orbx_map_add_points(m: *mut orbx_map, xyz: *const f64, desc: *const u8, has_desc: *const u8, observations: *const c_int, n: c_int, first_id: *mut i64) -> c_int;
    pub fn orbx_map_project(m: *mut orbx_map, R: *const f64, t: *const f64, xy: *mut f64, vis: *mut u8) -> c_int;
    pub fn orbx_map_find_matches(m: *mut orbx_map, matcher: *mut orbx_matcher, R: *const f64, t: *const f64, frame_desc: *const u8, n_frame: c_int, pairs: *mut c_int, cap: c_int, n_pairs: *mut c_int, n_visible: *mut c_int) -> c_int;
    pub fn orbx_map_update_observations(m: *mut orbx_map, pairs: *const c_int, n: c_int) -> c_int;
    pub fn orbx_map_prune_outliers(m: *mut orbx_map, removed: *mut c_int) -> c_int;
    pub fn orbx_map_stable_count(m: *mut orbx_map, count: *mut c_int) -> c_int;
    pub fn orbx_map_read_observations(m: *mut orbx_map, obs: *mut c_int, alive: *mut u8, n: c_int) -> c_int;
    pub fn orbx_pose_estimator_create(device: c_int, max_points: c_int, max_problems: c_int, fx: f64, fy: f64, cx: f64, cy: f64, out: *mut *mut orbx_pose_estimator) -> c_int;
    pub fn orbx_pose_estimator_destroy(pe: *mut orbx_pose_estimator);
    pub fn orbx_pose_estimator_set_stream(pe: *mut orbx_pose_estimator, cuda_stream: *mut c_void) -> c_int;
    pub fn orbx_pose_estimator_sync(pe: *mut orbx_pose_estimator) -> c_int;
    pub fn orbx_compute_essential_matrix(pe: *mut orbx_pose_estimator, pts1: *const f32, n1: c_int, pts2: *const f32, n2: c_int, E: *mut f64, mask: *mut u8, n_inliers: *mut c_int, iterations: *mut c_int) -> c_int;
    pub fn orbx_recover_pose(pe: *mut orbx_pose_estimator, E: *const f64, pts1: *const f32, n1: c_int, pts2: *const f32, n2: c_int, R: *mut f64, t: *mut f64, mask: *mut u8, n_good: *mut c_int) -> c_int;
    pub fn orbx_estimate_pose_batch(pe: *mut orbx_pose_estimator, pts1: *const f32, pts2: *const f32, n_dev: *const c_int, npairs: c_int, max_rows: c_int, E: *mut f64, R: *mut f64, t: *mut f64, n_inliers: *mut c_int, n_good: *mut c_int, status: *mut c_int, ransac_mask: *mut u8, pose_mask: *mut u8) -> c_int;
    pub fn orbx_triangulator_create(device: c_int, max_points: c_int, fx: f64, fy: f64, cx: f64, cy: f64, out: *mut *mut orbx_triangulator) -> c_int;
    pub fn orbx_triangulator_destroy(t: *mut orbx_triangulator);
    pub fn orbx_triangulator_set_thresholds(t: *mut orbx_triangulator, min_parallax_deg: f64, max_reproj_error: f64) -> c_int;
    pub fn orbx_triangulator_get_thresholds(t: *const orbx_triangulator, min_parallax_deg: *mut f64, max_reproj_error: *mut f64) -> c_int;
    pub fn orbx_triangulate(t: *mut orbx_triangulator, R1: *const f64, t1: *const f64, R2: *const f64, t2: *const f64, pts1: *const f32, n1: c_int, pts2: *const f32, n2: c_int, desc: *const u8, desc_rows: c_int, ids: *mut i32, xyz: *mut f64, out_desc: *mut u8, has_desc: *mut u8, points4d: *mut f32, cap: c_int, n_out: *mut c_int) -> c_int;
    pub fn orbx_triangulate_batch(t: *mut orbx_triangulator, R1: *const f64, t1: *const f64, R2: *const f64, t2: *const f64, pts1: *const f32, pts2: *const f32, n_dev: *const c_int, npairs: c_int, max_rows: c_int, ids: *mut i32, xyz: *mut f64, n_out: *mut c_int) -> c_int;
    pub fn orbx_triangulator_set_stream(t: *mut orbx_triangulator, cuda_stream: *mut c_void) -> c_int;
    pub fn orbx_transform_points_to_world(t: *mut orbx_triangulator, R: *const f64, tv: *const f64, xyz: *mut f64, n: c_int) -> c_int;
    pub fn orbx_sequence_create(device: c_int, width: c_int, height: c_int, nfeatures: c_int, max_frames: c_int, chunk: c_int, host_chunk: c_int, ratio: f32, cross_check: c_int, out: *mut *mut orbx_sequence) -> c_int;
    pub fn orbx_sequence_destroy(s: *mut orbx_sequence);
    pub fn orbx_sequence_capacity(s: *const orbx_sequence) -> c_int;
    pub fn orbx_sequence_set_reference_filter(s: *mut orbx_sequence, enable: c_int, ratio: f32) -> c_int;
    pub fn orbx_sequence_run_device(s: *mut orbx_sequence, frames_dev: *const u8, pitch: usize, frame_stride: usize, T: c_int) -> c_int;
    pub fn orbx_sequence_run_host(s: *mut orbx_sequence, frames: *const u8, T: c_int, kps: *mut orbx_keypoint, desc: *mut u8, n: *mut c_int, good: *mut orbx_dmatch, n_good: *mut c_int) -> c_int;
    pub fn orbx_sequence_wait_pass(s: *mut orbx_sequence, back: c_int) -> c_int;
    pub fn orbx_sequence_results(s: *mut orbx_sequence, kps: *mut *mut orbx_keypoint, desc: *mut *mut u8, n: *mut *mut c_int, good: *mut *mut orbx_dmatch, n_good: *mut *mut c_int) -> c_int;
    pub fn orbx_sequence_sync(s: *mut orbx_sequence) -> c_int;
    pub fn orbx_sequence_set_timing(s: *mut orbx_sequence, enable: c_int) -> c_int;
    pub fn orbx_sequence_elapsed(s: *mut orbx_sequence, ms: *mut f32, passes: *mut c_int, reset: c_int) -> c_int;
    pub fn orbx_comm_create(devices: *const c_int, ndev: c_int, out: *mut *mut orbx_comm) -> c_int;
    pub fn orbx_comm_destroy(c: *mut orbx_comm);
    pub fn orbx_comm_size(c: *const orbx_comm) -> c_int;
    pub fn orbx_comm_set_train(c: *mut orbx_comm, train: *const u8, nt: c_int) -> c_int;
    pub fn orbx_comm_knn_match2(c: *mut orbx_comm, q: *const u8, nq: c_int, out: *mut orbx_dmatch, n_per_q: *mut c_int) -> c_int;
    pub fn orbx_comm_match(c: *mut orbx_comm, q: *const u8, nq: c_int, out: *mut orbx_dmatch, n: *mut c_int) -> c_int;
    pub fn orbx_comm_sync(c: *mut orbx_comm) -> c_int;
    pub fn orbx_comm_set_timing(c: *mut orbx_comm, enable: c_int) -> c_int;
    pub fn orbx_comm_timing(c: *mut orbx_comm, ms: *mut f32, calls: *mut c_int, kernel_ms_rank0: *mut f32, reset: c_int) -> c_int;
    pub fn orbx_last_error() -> *const c_char;
    pub fn orbx_version() -> *const c_char;
    pub fn orbx_launch_count() -> u64;
    pub fn orbx_popc_microbench(device: c_int, mode: c_int, popc_per_s: *mut f64) -> c_int;
}
