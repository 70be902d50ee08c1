//! `gffx-cuda`: safe wrappers over the C ABI of include/gffx_b200.h for the three commands of GFFx's hot path
//! (src/commands/intersect.rs, depth.rs, coverage.rs of the reference).  Source a maintainer adds; not compiled in the
//! build image (no cargo / rustc there) — the same entry points are exercised through gffx_b200/api.py and the C++ CLI.
use anyhow::{anyhow, Result};
use std::{ffi::{c_char, c_int, c_void, CStr, CString}, ptr};

#[repr(C)] pub struct GffxIndex { _p: [u8; 0] }
#[repr(C)] pub struct GffxResult { _p: [u8; 0] }
#[repr(C)] pub struct GffxResultInfo {   // gffx_result_info, include/gffx_b200.h
    pub n_queries: u64, pub n_pairs: u64, pub n_roots: u64, pub n_ranges: u64, pub bytes_out: u64, pub n_warn: u64,
    pub pairs: *const u32, pub pairs_device: *const u32, pub ranges: *const u64, pub warn_fids: *const u32,
    pub kernel_ms: f64, pub stage_ms: [f64; 12], pub launches: u32, pub reserved: u32,
}
#[repr(C)] pub struct GffxCounts { _p: [u8; 0] }

#[repr(C)]
pub struct GffxIndexDesc {
    pub n_seqs: u32, pub seq_names: *const *const c_char,
    pub n_roots: u64, pub root_seq: *const u32, pub root_start: *const u32, pub root_end: *const u32, pub root_fid: *const u32,
    pub n_gof: u64, pub gof_fid: *const u32, pub gof_start: *const u64, pub gof_end: *const u64,
    pub gff_text: *const u8, pub gff_len: u64, pub device: c_int,
}

pub const F_DEVICE_INPUT: u32 = 1; pub const F_WANT_PAIRS: u32 = 2;

extern "C" {
    fn gffx_last_error() -> *const c_char;
    fn gffx_index_open(gff_path: *const c_char, device: c_int, out: *mut *mut GffxIndex) -> c_int;
    fn gffx_index_create(desc: *const GffxIndexDesc, out: *mut *mut GffxIndex) -> c_int;
    fn gffx_index_destroy(idx: *mut GffxIndex);
    fn gffx_query_features(idx: *mut GffxIndex, chr: *const u32, start: *const u32, end: *const u32, n: u64,
                           mode: c_int, invert: c_int, flags: u32, out: *mut *mut GffxResult) -> c_int;
    fn gffx_intersect(idx: *mut GffxIndex, chr: *const u32, start: *const u32, end: *const u32, n: u64, mode: c_int,
                      invert: c_int, entire_group: c_int, types_csv: *const c_char, flags: u32,
                      out: *mut *mut GffxResult) -> c_int;
    fn gffx_result_n_pairs(r: *const GffxResult) -> u64;
    fn gffx_result_pairs(r: *const GffxResult) -> *const u32;
    fn gffx_result_n_ranges(r: *const GffxResult) -> u64;
    fn gffx_result_ranges(r: *const GffxResult) -> *const u64;
    fn gffx_result_n_warn(r: *const GffxResult) -> u64;
    fn gffx_result_warn_fids(r: *const GffxResult) -> *const u32;
    // the whole summary in one call (#[repr(C)] struct GffxResultInfo mirrors gffx_result_info field by field)
    fn gffx_result_get_info(r: *const GffxResult, out: *mut GffxResultInfo) -> c_int;
    fn gffx_result_free(r: *mut GffxResult);
    // flags: 1 DEVICE_INPUT, 2 WANT_PAIRS, 4 NO_HOST_COPY, 8 PAIRS_ON_DEVICE, 16 STAGE_TIMES (per-stage device ms in the result:
    // a dozen CUDA events inside the call, ~40 us; leave it off outside profiling)
    fn gffx_write_ranges(idx: *const GffxIndex, ranges: *const u64, n: u64, out_path: *const c_char) -> c_int;
    fn gffx_parse_bed_device(idx: *mut GffxIndex, bed_path: *const c_char, dialect: c_int, d_chr: *mut *const u32, d_start: *mut *const u32,
                             d_end: *mut *const u32, n: *mut u64) -> c_int;   // regions stay in HBM: pass them on with GFFX_F_DEVICE_INPUT (flag 1)
    fn gffx_gather_text(idx: *mut GffxIndex, ranges: *const u64, n: u64, flags: u32, text: *mut *const u8, n_bytes: *mut u64) -> c_int;
    fn gffx_write_text(text: *const u8, n_bytes: u64, out_path: *const c_char) -> c_int;
    fn gffx_parse_reads(idx: *const GffxIndex, path: *const c_char, chr: *mut *mut u32, start: *mut *mut u32, end: *mut *mut u32, n: *mut u64,
                        stats: *mut u64) -> c_int;
    fn gffx_index_lines_info(idx: *const GffxIndex, digest: *mut u64, source: *mut c_int) -> c_int;
    fn gffx_free(p: *mut std::ffi::c_void);
    // allocation-free join-pair count of a device-resident batch (no counterpart in the reference; include/gffx_b200.h)
    pub fn gffx_count_pairs_noalloc(idx: *mut GffxIndex, d_chr: *const u32, d_start: *const u32, d_end: *const u32, n: u64, mode: c_int, invert: c_int,
                                    d_out: *mut u64, stream: *mut std::ffi::c_void) -> c_int;
    fn gffx_counts_begin(idx: *mut GffxIndex, kind: c_int, out: *mut *mut GffxCounts) -> c_int;
    fn gffx_counts_add(st: *mut GffxCounts, chr: *const u32, start: *const u32, end: *const u32, n: u64, flags: u32) -> c_int;
    fn gffx_counts_write(st: *mut GffxCounts, out_path: *const c_char) -> c_int;
    fn gffx_counts_free(st: *mut GffxCounts);
}

fn check(rc: c_int) -> Result<()> {
    if rc == 0 { return Ok(()); }
    let msg = unsafe { CStr::from_ptr(gffx_last_error()) }.to_string_lossy().into_owned();
    Err(anyhow!("gffx-cuda [{rc}]: {msg}"))
}

pub struct DeviceIndex(*mut GffxIndex);
unsafe impl Send for DeviceIndex {}
impl Drop for DeviceIndex { fn drop(&mut self) { unsafe { gffx_index_destroy(self.0) } } }

impl DeviceIndex {
    /// replaces TreeIndexData::load_tree_index + load_gof + the GFF mmap
    pub fn open(gff: &std::path::Path, device: i32) -> Result<Self> {
        let p = CString::new(gff.to_string_lossy().as_bytes())?;
        let mut h = ptr::null_mut();
        check(unsafe { gffx_index_open(p.as_ptr(), device, &mut h) })?;
        Ok(Self(h))
    }

    /// drop-in for `query_features(&index, &regions, mode, invert)` (intersect.rs:105-169)
    pub fn query_features(&self, regions: &[(u32, u32, u32)], mode: i32, invert: bool) -> Result<Vec<(u32, u32, u32)>> {
        let (c, s, e): (Vec<u32>, Vec<u32>, Vec<u32>) = unzip3(regions);
        let mut r = ptr::null_mut();
        check(unsafe { gffx_query_features(self.0, c.as_ptr(), s.as_ptr(), e.as_ptr(), c.len() as u64, mode,
                                           invert as c_int, F_WANT_PAIRS, &mut r) })?;
        let n = unsafe { gffx_result_n_pairs(r) } as usize;
        let p = unsafe { std::slice::from_raw_parts(gffx_result_pairs(r), 3 * n) };
        let v = p.chunks_exact(3).map(|t| (t[0], t[1], t[2])).collect();
        unsafe { gffx_result_free(r) };
        Ok(v)
    }

    /// the read source of depth / coverage from a SAM / BAM file: what process_bam (depth.rs:297-423) and
    /// collect_by_root_from_bam (coverage.rs:125-205) take from rust-htslib, as (chr_id, start, end) triples in file order
    pub fn parse_reads(&self, path: &std::path::Path) -> Result<Vec<(u32, u32, u32)>> {
        let p = CString::new(path.to_string_lossy().as_bytes())?;
        let (mut c, mut s, mut e, mut n) = (ptr::null_mut(), ptr::null_mut(), ptr::null_mut(), 0u64);
        check(unsafe { gffx_parse_reads(self.0, p.as_ptr(), &mut c, &mut s, &mut e, &mut n, ptr::null_mut()) })?;
        let n = n as usize;
        let v = unsafe {
            let (cs, ss, es) = (std::slice::from_raw_parts(c, n), std::slice::from_raw_parts(s, n), std::slice::from_raw_parts(e, n));
            (0..n).map(|i| (cs[i], ss[i], es[i])).collect()
        };
        unsafe { gffx_free(c as *mut _); gffx_free(s as *mut _); gffx_free(e as *mut _) };
        Ok(v)
    }

    /// drop-in for the body of intersect::run after parse_bed_file (intersect.rs:586-654)
    pub fn intersect_to(&self, regions: &[(u32, u32, u32)], mode: i32, invert: bool, entire_group: bool,
                        types: Option<&str>, out: Option<&std::path::Path>) -> Result<()> {
        let (c, s, e): (Vec<u32>, Vec<u32>, Vec<u32>) = unzip3(regions);
        let t = types.map(CString::new).transpose()?;
        let mut r = ptr::null_mut();
        check(unsafe { gffx_intersect(self.0, c.as_ptr(), s.as_ptr(), e.as_ptr(), c.len() as u64, mode, invert as c_int,
                                      entire_group as c_int, t.as_ref().map_or(ptr::null(), |x| x.as_ptr()), 0, &mut r) })?;
        for k in 0..unsafe { gffx_result_n_warn(r) } as usize {   // the reference's stderr text, intersect.rs:269-272
            eprintln!("[WARN] skipped fid={}", unsafe { *gffx_result_warn_fids(r).add(k) });
        }
        let o = out.map(|p| CString::new(p.to_string_lossy().as_bytes())).transpose()?;
        // record emission on the device (K8): the ranges are still in HBM, the GFF text is resident there too; one
        // gathered buffer comes back and goes out with one write.  (gffx_write_ranges is the host-side writev path.)
        let (mut text, mut nb) = (ptr::null(), 0u64);
        let rc = unsafe { gffx_gather_text(self.0, ptr::null(), gffx_result_n_ranges(r), 0, &mut text, &mut nb) };
        unsafe { gffx_result_free(r) };
        check(rc)?;
        check(unsafe { gffx_write_text(text, nb, o.as_ref().map_or(ptr::null(), |x| x.as_ptr())) })
    }
}
fn unzip3(v: &[(u32, u32, u32)]) -> (Vec<u32>, Vec<u32>, Vec<u32>) {
    let mut a = Vec::with_capacity(v.len()); let mut b = a.clone(); let mut c = a.clone();
    for &(x, y, z) in v { a.push(x); b.push(y); c.push(z); }
    (a, b, c)
}
