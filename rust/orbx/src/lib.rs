//! Safe mirror of slamkit-rs `src/feature/{detector,matcher}.rs`, `odometry::PoseEstimator` and `mapping::Triangulator` over liborbx.
//! Same method names and error behaviour: working calls take `&mut self`, empty inputs are not errors,
//! results are freshly allocated.  NOT compiled in this repository's build image (no Rust toolchain).
use orbx_sys as sys;
use std::{error::Error, ffi::CStr, ptr};

pub type KeyPoint = sys::orbx_keypoint; // layout == cv::KeyPoint
pub type DMatch = sys::orbx_dmatch;     // layout == cv::DMatch

fn err(code: i32) -> Box<dyn Error> {
    let msg = unsafe { CStr::from_ptr(sys::orbx_last_error()) }.to_string_lossy().into_owned();
    format!("liborbx error {code}: {msg}").into()
}

/// A borrowed 8-bit single-channel image (what `&Mat` CV_8UC1 is on the reference side).
pub struct Gray<'a> { pub data: &'a [u8], pub width: i32, pub height: i32, pub pitch: usize }

pub struct OrbDetector { h: *mut sys::orbx_detector, max_features: i32, nfeatures: i32, device: i32, max_w: i32, max_h: i32, graph: bool }

impl OrbDetector {
    /// detector.rs:15-18 - unlike the reference (which forwards to `ORB::create_def()` and keeps 500), `max_features`
    /// is honoured; `new_reference_faithful` restores the reference's behaviour.  Device scratch is sized on the first
    /// image (and re-sized when a larger image arrives), not for the 4095 x 4095 maximum.
    pub fn new(max_features: i32) -> Result<Self, Box<dyn Error>> { Self::with_options(max_features, max_features, 0) }
    /// `ORB::create_def()` exactly as detector.rs:16: nfeatures stays 500, `max_features` is only echoed (detector.rs:42-44).
    pub fn new_reference_faithful(max_features: i32) -> Result<Self, Box<dyn Error>> { Self::with_options(max_features, 500, 0) }
    pub fn with_options(max_features: i32, nfeatures: i32, device: i32) -> Result<Self, Box<dyn Error>> {
        Ok(Self { h: ptr::null_mut(), max_features, nfeatures, device, max_w: 0, max_h: 0, graph: false })
    }
    /// Replay single-frame calls as one CUDA graph (orbx_detector_set_graph): same bytes, lower per-call latency.
    pub fn set_graph(&mut self, enable: bool) { self.graph = enable; if !self.h.is_null() { unsafe { sys::orbx_detector_set_graph(self.h, enable as i32) }; } }
    fn ensure(&mut self, w: i32, h: i32) -> Result<(), Box<dyn Error>> {
        if !self.h.is_null() && w <= self.max_w && h <= self.max_h { return Ok(()); }
        if !self.h.is_null() { unsafe { sys::orbx_detector_destroy(self.h) }; self.h = ptr::null_mut(); }
        let mut p = unsafe { let mut p = std::mem::zeroed(); sys::orbx_default_params(&mut p); p };
        p.nfeatures = self.nfeatures;
        let (mw, mh) = (w.max(self.max_w).max(16), h.max(self.max_h).max(16));
        let rc = unsafe { sys::orbx_detector_create(&p, self.device, mw, mh, 1, &mut self.h) };
        if rc != 0 { self.h = ptr::null_mut(); return Err(err(rc)); }
        self.max_w = mw; self.max_h = mh;
        if self.graph { unsafe { sys::orbx_detector_set_graph(self.h, 1) }; }
        Ok(())
    }
    fn run(&mut self, img: &Gray, with_desc: bool) -> Result<(Vec<KeyPoint>, Vec<u8>), Box<dyn Error>> {
        self.ensure(img.width, img.height)?;
        let mut cap = (self.nfeatures + self.nfeatures / 4 + 256) as usize;
        loop {
            let mut kps = vec![KeyPoint::default(); cap];
            let mut desc = vec![0u8; if with_desc { cap * 32 } else { 0 }];
            let mut n = 0i32;
            let rc = unsafe {
                if with_desc { sys::orbx_detect_and_compute(self.h, img.data.as_ptr(), img.width, img.height, img.pitch, kps.as_mut_ptr(), desc.as_mut_ptr(), cap as i32, &mut n) }
                else { sys::orbx_detect(self.h, img.data.as_ptr(), img.width, img.height, img.pitch, kps.as_mut_ptr(), cap as i32, &mut n) }
            };
            if rc == sys::ORBX_E_CAPACITY && n as usize > cap { cap = n as usize; continue; }
            if rc != 0 { return Err(err(rc)); }
            kps.truncate(n as usize); desc.truncate(n as usize * 32);
            return Ok((kps, desc));
        }
    }
    /// detector.rs:20-24
    pub fn detect(&mut self, image: &Gray) -> Result<Vec<KeyPoint>, Box<dyn Error>> { Ok(self.run(image, false)?.0) }
    /// detector.rs:26-40 — descriptors are K x 32 bytes, row-major (== Mat K x 32 CV_8U)
    pub fn detect_and_compute(&mut self, image: &Gray) -> Result<(Vec<KeyPoint>, Vec<u8>), Box<dyn Error>> { self.run(image, true) }
    /// detector.rs:42-44
    pub fn max_features(&self) -> i32 { self.max_features }
}
impl Drop for OrbDetector { fn drop(&mut self) { if !self.h.is_null() { unsafe { sys::orbx_detector_destroy(self.h) } } } }

pub struct FeatureMatcher { h: *mut sys::orbx_matcher }

impl FeatureMatcher {
    /// matcher.rs:15-21: `BFMatcher::create(NORM_HAMMING, crossCheck = false)`
    pub fn new() -> Result<Self, Box<dyn Error>> { Self::with_options(false, 0) }
    /// `BFMatcher::create(NORM_HAMMING, crossCheck = true)`: `match_descriptors` keeps mutual nearest pairs only, and
    /// `knn_match` zeroes the count of queries that fail the mutual test (the north_star's k=2 + cross-check composition)
    pub fn new_cross_check() -> Result<Self, Box<dyn Error>> { Self::with_options(true, 0) }
    pub fn with_options(cross_check: bool, device: i32) -> Result<Self, Box<dyn Error>> {
        let mut h = ptr::null_mut();
        let rc = unsafe { sys::orbx_matcher_create(device, cross_check as i32, &mut h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(Self { h })
    }
    /// `BFMatcher::knnMatch(k = 2)`: per query row up to two matches in ascending (distance, train index) order
    pub fn knn_match(&mut self, desc1: &[u8], desc2: &[u8]) -> Result<Vec<Vec<DMatch>>, Box<dyn Error>> {
        let (nq, nt) = (desc1.len() / 32, desc2.len() / 32);
        if nq == 0 { return Ok(Vec::new()); }
        let mut out = vec![DMatch::default(); nq * 2];
        let mut cnt = vec![0i32; nq];
        let rc = unsafe { sys::orbx_knn_match2(self.h, desc1.as_ptr(), nq as i32, desc2.as_ptr(), nt as i32, out.as_mut_ptr(), cnt.as_mut_ptr()) };
        if rc != 0 { return Err(err(rc)); }
        Ok((0..nq).map(|i| out[2 * i..2 * i + (cnt[i].clamp(0, 2) as usize)].to_vec()).collect())
    }
    /// Lowe's ratio test on a `knn_match` result: keeps m[0] iff two matches exist and m[0].distance < ratio * m[1].distance
    pub fn ratio_test(&self, knn: &[Vec<DMatch>], ratio: f32) -> Vec<DMatch> {
        knn.iter().filter(|m| m.len() == 2 && m[0].distance < ratio * m[1].distance).map(|m| m[0]).collect()
    }
    /// matcher.rs:24-37 — desc1/desc2: N x 32 bytes; either side empty -> empty result
    pub fn match_descriptors(&mut self, desc1: &[u8], desc2: &[u8]) -> Result<Vec<DMatch>, Box<dyn Error>> {
        let (nq, nt) = (desc1.len() / 32, desc2.len() / 32);
        if nq == 0 || nt == 0 { return Ok(Vec::new()); }
        let mut out = vec![DMatch::default(); nq];
        let mut n = 0i32;
        let rc = unsafe { sys::orbx_match(self.h, desc1.as_ptr(), nq as i32, desc2.as_ptr(), nt as i32, out.as_mut_ptr(), &mut n) };
        if rc != 0 { return Err(err(rc)); }
        out.truncate(n as usize);
        Ok(out)
    }
    /// matcher.rs:40-66
    pub fn filter_good_matches(&self, matches: &[DMatch], ratio: f32) -> Vec<DMatch> {
        if matches.is_empty() { return Vec::new(); }
        let mut out = vec![DMatch::default(); matches.len()];
        let mut n = 0i32;
        let rc = unsafe { sys::orbx_filter_good_matches(self.h, matches.as_ptr(), matches.len() as i32, ratio, out.as_mut_ptr(), &mut n) };
        if rc != 0 { return Vec::new(); }
        out.truncate(n as usize);
        out
    }
    /// odometry/pose.rs:27-44 (`PoseEstimator::extract_matched_points`): (points1, points2) as (x, y) pairs;
    /// `Err` when a match index is outside its keypoint slice, like `Vector::get(..)?` in the reference.
    pub fn extract_matched_points(&self, kp1: &[KeyPoint], kp2: &[KeyPoint], matches: &[DMatch]) -> Result<(Vec<[f32; 2]>, Vec<[f32; 2]>), Box<dyn Error>> {
        let mut p1 = vec![[0f32; 2]; matches.len()];
        let mut p2 = vec![[0f32; 2]; matches.len()];
        if matches.is_empty() { return Ok((p1, p2)); }
        let rc = unsafe { sys::orbx_extract_matched_points(self.h, kp1.as_ptr(), kp1.len() as i32, kp2.as_ptr(), kp2.len() as i32, matches.as_ptr(), matches.len() as i32, p1.as_mut_ptr() as *mut f32, p2.as_mut_ptr() as *mut f32) };
        if rc != 0 { return Err(err(rc)); }
        Ok((p1, p2))
    }
}
impl Drop for FeatureMatcher { fn drop(&mut self) { unsafe { sys::orbx_matcher_destroy(self.h) } } }

/// odometry::CameraIntrinsics (src/odometry/camera.rs:5-36)
#[derive(Clone, Copy, Debug)]
pub struct CameraIntrinsics { pub fx: f64, pub fy: f64, pub cx: f64, pub cy: f64 }
impl CameraIntrinsics {
    pub fn new(fx: f64, fy: f64, cx: f64, cy: f64) -> Self { Self { fx, fy, cx, cy } }
    pub fn kitti() -> Self { Self { fx: 718.856, fy: 718.856, cx: 607.1928, cy: 185.2157 } }
    /// camera.rs:29-36
    pub fn webcam_vga() -> Self { Self { fx: 500.0, fy: 500.0, cx: 320.0, cy: 240.0 } }
    /// camera.rs:39-54 (row-major 3x3; behind the `opencv-types` feature this becomes a CV_64F `Mat`)
    pub fn to_matrix(&self) -> [f64; 9] { [self.fx, 0.0, self.cx, 0.0, self.fy, self.cy, 0.0, 0.0, 1.0] }
}

/// (R row-major, t): world-to-camera pose as the ABI takes it (nalgebra's `Matrix3` is column-major: pass `r.transpose()`).
pub type Pose = ([f64; 9], [f64; 3]);

/// Mirror of odometry::PoseEstimator (src/odometry/pose.rs:10-114): cv::findEssentialMat(RANSAC, 0.999, 1.0, 1000) and
/// cv::recoverPose reproduced on the device with OpenCV's own RANSAC draws.
pub struct PoseEstimator { h: *mut sys::orbx_pose_estimator, matcher: FeatureMatcher }

impl PoseEstimator {
    /// pose.rs:17-23
    pub fn new(intrinsics: CameraIntrinsics) -> Result<Self, Box<dyn Error>> {
        let mut h = ptr::null_mut();
        let rc = unsafe { sys::orbx_pose_estimator_create(0, 1 << 14, 1, intrinsics.fx, intrinsics.fy, intrinsics.cx, intrinsics.cy, &mut h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(Self { h, matcher: FeatureMatcher::new()? })
    }
    /// pose.rs:27-44
    pub fn extract_matched_points(&self, kp1: &[KeyPoint], kp2: &[KeyPoint], matches: &[DMatch]) -> Result<(Vec<[f32; 2]>, Vec<[f32; 2]>), Box<dyn Error>> {
        self.matcher.extract_matched_points(kp1, kp2, matches)
    }
    /// pose.rs:47-80 — E row-major; `Err("Insufficient points: n (need 8)")` / `Err("Failed to compute essential matrix")`
    pub fn compute_essential_matrix(&self, points1: &[[f32; 2]], points2: &[[f32; 2]]) -> Result<[f64; 9], Box<dyn Error>> {
        let mut e = [0f64; 9];
        let rc = unsafe { sys::orbx_compute_essential_matrix(self.h, points1.as_ptr() as *const f32, points1.len() as i32,
            points2.as_ptr() as *const f32, points2.len() as i32, e.as_mut_ptr(), ptr::null_mut(), ptr::null_mut(), ptr::null_mut()) };
        if rc != 0 { return Err(err(rc)); }
        Ok(e)
    }
    /// pose.rs:83-114 — `Err("Too few inliers: k")` when fewer than 8 points are in front of both cameras
    pub fn recover_pose(&self, essential: &[f64; 9], points1: &[[f32; 2]], points2: &[[f32; 2]]) -> Result<Pose, Box<dyn Error>> {
        let (mut r, mut t) = ([0f64; 9], [0f64; 3]);
        let rc = unsafe { sys::orbx_recover_pose(self.h, essential.as_ptr(), points1.as_ptr() as *const f32, points1.len() as i32,
            points2.as_ptr() as *const f32, points2.len() as i32, r.as_mut_ptr(), t.as_mut_ptr(), ptr::null_mut(), ptr::null_mut()) };
        if rc != 0 { return Err(err(rc)); }
        Ok((r, t))
    }
}
impl Drop for PoseEstimator { fn drop(&mut self) { unsafe { sys::orbx_pose_estimator_destroy(self.h) } } }

/// mapping::MapPoint (src/mapping/triangulation.rs:11-22)
#[derive(Clone, Debug)]
pub struct MapPoint { pub position: [f64; 3], pub descriptor: Option<Vec<u8>>, pub observations: usize, pub id: usize }

/// Mirror of mapping::Triangulator (src/mapping/triangulation.rs:52-198)
pub struct Triangulator { h: *mut sys::orbx_triangulator }

impl Triangulator {
    /// triangulation.rs:62-68
    pub fn new(intrinsics: CameraIntrinsics) -> Result<Self, Box<dyn Error>> {
        let mut h = ptr::null_mut();
        let rc = unsafe { sys::orbx_triangulator_create(0, 1 << 16, intrinsics.fx, intrinsics.fy, intrinsics.cx, intrinsics.cy, &mut h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(Self { h })
    }
    fn thresholds(&self) -> (f64, f64) {
        let (mut a, mut b) = (0f64, 0f64);
        unsafe { sys::orbx_triangulator_get_thresholds(self.h, &mut a, &mut b) };
        (a, b)
    }
    /// triangulation.rs:71-74
    pub fn with_min_parallax(self, deg: f64) -> Self { unsafe { sys::orbx_triangulator_set_thresholds(self.h, deg, self.thresholds().1) }; self }
    /// triangulation.rs:77-80
    pub fn with_max_reproj_error(self, error: f64) -> Self { unsafe { sys::orbx_triangulator_set_thresholds(self.h, self.thresholds().0, error) }; self }
    /// triangulation.rs:92-158 — `descriptors`: rows x 32 bytes (row i goes to the point of correspondence i)
    pub fn triangulate(&self, pose1: &Pose, pose2: &Pose, points1: &[[f32; 2]], points2: &[[f32; 2]], descriptors: Option<&[u8]>) -> Result<Vec<MapPoint>, Box<dyn Error>> {
        let cap = points1.len().max(1);
        let (mut ids, mut xyz) = (vec![0i32; cap], vec![0f64; 3 * cap]);
        let (mut desc, mut has) = (vec![0u8; 32 * cap], vec![0u8; cap]);
        let mut m = 0i32;
        let (dptr, drows) = descriptors.map_or((ptr::null(), 0), |d| (d.as_ptr(), (d.len() / 32) as i32));
        let rc = unsafe { sys::orbx_triangulate(self.h, pose1.0.as_ptr(), pose1.1.as_ptr(), pose2.0.as_ptr(), pose2.1.as_ptr(),
            points1.as_ptr() as *const f32, points1.len() as i32, points2.as_ptr() as *const f32, points2.len() as i32, dptr, drows,
            ids.as_mut_ptr(), xyz.as_mut_ptr(), desc.as_mut_ptr(), has.as_mut_ptr(), ptr::null_mut(), cap as i32, &mut m) };
        if rc != 0 { return Err(err(rc)); }
        Ok((0..m as usize).map(|k| MapPoint { position: [xyz[3 * k], xyz[3 * k + 1], xyz[3 * k + 2]], id: ids[k] as usize, observations: 1,
            descriptor: (has[k] != 0).then(|| desc[32 * k..32 * k + 32].to_vec()) }).collect())
    }
}
impl Drop for Triangulator { fn drop(&mut self) { unsafe { sys::orbx_triangulator_destroy(self.h) } } }

/// One frame's features and (for t >= 1) the good matches of the pair (t-1, t), as `Sequence::run` returns them.
pub struct FrameResult { pub keypoints: Vec<KeyPoint>, pub descriptors: Vec<u8>, pub good_matches: Vec<DMatch> }

/// The loop of examples/visual_odometry.rs:84-218 (detect_and_compute :102, match_descriptors :106, filter :107) over a
/// whole sequence in one call: liborbx schedules the frames through its batched kernels (csrc/sequence.cu), upload,
/// kernels and download of consecutive chunks overlapping.  Matching is BFMatcher k=2 + cross-check + Lowe ratio.
pub struct Sequence { h: *mut sys::orbx_sequence, width: usize, height: usize }

impl Sequence {
    pub fn new(width: usize, height: usize, max_features: i32, max_frames: usize) -> Result<Self, Box<dyn Error>> {
        let mut h = ptr::null_mut();
        let rc = unsafe { sys::orbx_sequence_create(0, width as i32, height as i32, max_features, max_frames as i32, 1000, 250, 0.75, 1, &mut h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(Self { h, width, height })
    }
    /// Match the pairs exactly as the reference's loop does (visual_odometry.rs:106-107): `match_descriptors` (k=1, no
    /// cross-check) then `filter_good_matches(ratio)` (2.0 there) instead of k=2 + cross-check + Lowe.
    pub fn with_reference_filter(self, ratio: f32) -> Self { unsafe { sys::orbx_sequence_set_reference_filter(self.h, 1, ratio) }; self }
    /// `frames`: T tightly packed 8-bit gray images (T * height * width bytes).  A frame with more keypoints than the row
    /// capacity (a tie flood) is truncated to it and reported as an `Err` by the sync below.
    pub fn run(&mut self, frames: &[u8]) -> Result<Vec<FrameResult>, Box<dyn Error>> {
        let t = frames.len() / (self.width * self.height);
        let cap = unsafe { sys::orbx_sequence_capacity(self.h) } as usize;
        let mut kps = vec![KeyPoint::default(); t * cap];
        let mut desc = vec![0u8; t * cap * 32];
        let mut good = vec![DMatch::default(); t * cap];
        let (mut n, mut n_good) = (vec![0i32; t], vec![0i32; t]);
        let rc = unsafe { sys::orbx_sequence_run_host(self.h, frames.as_ptr(), t as i32, kps.as_mut_ptr(), desc.as_mut_ptr(),
            n.as_mut_ptr(), good.as_mut_ptr(), n_good.as_mut_ptr()) };
        if rc != 0 { return Err(err(rc)); }
        let rc = unsafe { sys::orbx_sequence_sync(self.h) };
        if rc != 0 { return Err(err(rc)); }
        Ok((0..t).map(|i| {
            let (k, g) = ((n[i].max(0) as usize).min(cap), if i > 0 { (n_good[i].max(0) as usize).min(cap) } else { 0 });
            FrameResult { keypoints: kps[i * cap..i * cap + k].to_vec(), descriptors: desc[i * cap * 32..(i * cap + k) * 32].to_vec(),
                          good_matches: good[i * cap..i * cap + g].to_vec() }
        }).collect())
    }
}
impl Sequence {
    /// Streaming form of `run`: queue a sequence into caller-owned result arrays (`cap` = `capacity()` rows per frame) and
    /// return at once; consecutive calls are pipelined into each other.  The arrays belong to the library until
    /// `wait_pass` / `sync` says the call has finished, hence `unsafe`.
    pub unsafe fn run_async(&mut self, frames: &[u8], kps: *mut KeyPoint, desc: *mut u8, n: *mut i32, good: *mut DMatch,
                            n_good: *mut i32) -> Result<(), Box<dyn Error>> {
        let t = frames.len() / (self.width * self.height);
        let rc = sys::orbx_sequence_run_host(self.h, frames.as_ptr(), t as i32, kps, desc, n, good, n_good);
        if rc != 0 { return Err(err(rc)); }
        Ok(())
    }
    pub fn capacity(&self) -> usize { unsafe { sys::orbx_sequence_capacity(self.h) as usize } }
    /// Block until the call issued `back` calls before the most recent one has finished (0 = the most recent).
    pub fn wait_pass(&mut self, back: i32) -> Result<(), Box<dyn Error>> {
        let rc = unsafe { sys::orbx_sequence_wait_pass(self.h, back) };
        if rc != 0 { return Err(err(rc)); }
        Ok(())
    }
    pub fn sync(&mut self) -> Result<(), Box<dyn Error>> {
        let rc = unsafe { sys::orbx_sequence_sync(self.h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(())
    }
}
impl Drop for Sequence { fn drop(&mut self) { unsafe { sys::orbx_sequence_destroy(self.h) } } }

/// Mirror of mapping::Map (src/mapping/map.rs:8-150) with positions, descriptors and observation counts resident in
/// HBM (orbx_map_*): ids are handed out sequentially as in map.rs:31-37.
pub struct Map { h: *mut sys::orbx_map }

impl Map {
    /// map.rs:21-28; `capacity` = total number of points ever added
    pub fn new(intrinsics: CameraIntrinsics, capacity: usize) -> Result<Self, Box<dyn Error>> {
        let mut h = ptr::null_mut();
        let rc = unsafe { sys::orbx_map_create(0, capacity as i32, intrinsics.fx, intrinsics.fy, intrinsics.cx, intrinsics.cy, &mut h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(Self { h })
    }
    /// map.rs:31-37: returns the id of the first new point
    pub fn add_points(&mut self, points: &[MapPoint]) -> Result<usize, Box<dyn Error>> {
        let n = points.len();
        let mut xyz = Vec::with_capacity(3 * n);
        let (mut desc, mut has, mut obs) = (vec![0u8; 32 * n], vec![0u8; n], vec![0i32; n]);
        for (i, p) in points.iter().enumerate() {
            xyz.extend_from_slice(&p.position);
            if let Some(d) = &p.descriptor { desc[32 * i..32 * i + 32].copy_from_slice(&d[..32]); has[i] = 1; }
            obs[i] = p.observations as i32;
        }
        let mut first = 0i64;
        let rc = unsafe { sys::orbx_map_add_points(self.h, xyz.as_ptr(), desc.as_ptr(), has.as_ptr(), obs.as_ptr(), n as i32, &mut first) };
        if rc != 0 { return Err(err(rc)); }
        Ok(first as usize)
    }
    /// map.rs:45-47
    pub fn size(&self) -> usize { unsafe { sys::orbx_map_size(self.h) as usize } }
    /// map.rs:72-118: projection cull, `match_descriptors(map, frame)` (query = map rows) and `filter_good_matches(2.0)`;
    /// returns (map_point_id, keypoint_idx) pairs.  `descriptors`: K x 32 bytes of the frame.
    pub fn find_matches(&mut self, descriptors: &[u8], pose: &Pose, matcher: &mut FeatureMatcher) -> Result<Vec<(usize, usize)>, Box<dyn Error>> {
        let cap = self.size().max(1);
        let mut pairs = vec![0i32; 2 * cap];
        let (mut n, mut vis) = (0i32, 0i32);
        let rc = unsafe { sys::orbx_map_find_matches(self.h, matcher.h, pose.0.as_ptr(), pose.1.as_ptr(), descriptors.as_ptr(),
            (descriptors.len() / 32) as i32, pairs.as_mut_ptr(), cap as i32, &mut n, &mut vis) };
        if rc != 0 { return Err(err(rc)); }
        Ok((0..n as usize).map(|i| (pairs[2 * i] as usize, pairs[2 * i + 1] as usize)).collect())
    }
    /// map.rs:121-127
    pub fn update_observations(&mut self, matches: &[(usize, usize)]) {
        let flat: Vec<i32> = matches.iter().flat_map(|&(a, b)| [a as i32, b as i32]).collect();
        unsafe { sys::orbx_map_update_observations(self.h, flat.as_ptr(), matches.len() as i32) };
    }
    /// map.rs:130-137
    pub fn prune_outliers(&mut self) -> usize { let mut r = 0i32; unsafe { sys::orbx_map_prune_outliers(self.h, &mut r) }; r as usize }
    /// number of map.rs:140-145 `stable_points()`
    pub fn stable_count(&self) -> usize { let mut c = 0i32; unsafe { sys::orbx_map_stable_count(self.h, &mut c) }; c as usize }
    /// map.rs:147-150
    pub fn clear(&mut self) { unsafe { sys::orbx_map_clear(self.h) }; }
}
impl Drop for Map { fn drop(&mut self) { unsafe { sys::orbx_map_destroy(self.h) } } }

/// Query-sharded map-to-frame matching over the GPUs of one box (orbx_comm_*, csrc/comm.cu): the train set (the map's
/// descriptors, map.rs:107) is made resident on every GPU with one NCCL broadcast, the query rows are sliced over the
/// GPUs and every GPU's kernel stores its records straight into GPU `devices[0]`'s buffer.
pub struct ShardedMatcher { h: *mut sys::orbx_comm }

impl ShardedMatcher {
    pub fn new(devices: &[i32]) -> Result<Self, Box<dyn Error>> {
        let mut h = ptr::null_mut();
        let rc = unsafe { sys::orbx_comm_create(devices.as_ptr(), devices.len() as i32, &mut h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(Self { h })
    }
    /// `train`: T x 32 bytes (host); call again after every map update
    pub fn set_train(&mut self, train: &[u8]) -> Result<(), Box<dyn Error>> {
        let rc = unsafe { sys::orbx_comm_set_train(self.h, train.as_ptr(), (train.len() / 32) as i32) };
        if rc != 0 { return Err(err(rc)); }
        let rc = unsafe { sys::orbx_comm_sync(self.h) };
        if rc != 0 { return Err(err(rc)); }
        Ok(())
    }
    /// matcher.rs:24-37 over all GPUs: one DMatch per query row
    pub fn match_descriptors(&mut self, desc: &[u8]) -> Result<Vec<DMatch>, Box<dyn Error>> {
        let nq = desc.len() / 32;
        let mut out = vec![DMatch::default(); nq.max(1)];
        let mut n = 0i32;
        let rc = unsafe { sys::orbx_comm_match(self.h, desc.as_ptr(), nq as i32, out.as_mut_ptr(), &mut n) };
        if rc != 0 { return Err(err(rc)); }
        out.truncate(n as usize);
        Ok(out)
    }
    /// `BFMatcher::knnMatch(k = 2)` over all GPUs
    pub fn knn_match(&mut self, desc: &[u8]) -> Result<Vec<Vec<DMatch>>, Box<dyn Error>> {
        let nq = desc.len() / 32;
        if nq == 0 { return Ok(Vec::new()); }
        let mut out = vec![DMatch::default(); nq * 2];
        let mut cnt = vec![0i32; nq];
        let rc = unsafe { sys::orbx_comm_knn_match2(self.h, desc.as_ptr(), nq as i32, out.as_mut_ptr(), cnt.as_mut_ptr()) };
        if rc != 0 { return Err(err(rc)); }
        Ok((0..nq).map(|i| out[2 * i..2 * i + (cnt[i].clamp(0, 2) as usize)].to_vec()).collect())
    }
}
impl Drop for ShardedMatcher { fn drop(&mut self) { unsafe { sys::orbx_comm_destroy(self.h) } } }
