//! `GpuIndex`: the reference's `Index<Sketch>` (src/index/mod.rs:21-43) served by libhnswb200 on a B200.
//! UNCOMPILED (no Rust toolchain in the build image); see rust/README.md.
//!
//! `Sketch` is `struct Sketch { data: [u64; 16] }` (cli/src/sketch.rs:7-10); mark it `#[repr(C)]` so that `&[Sketch]`
//! is a contiguous `n x 128 B` buffer and the pointer casts below are sound -- no repacking on the way in.
use std::ffi::{CStr, CString};
use std::path::Path;

use hnsw_b200_sys as sys;
use hnsw_itu::{Distance, Index};
use hnsw_itu_cli::Sketch;

fn check(rc: i32) -> anyhow::Result<()> {
    if rc == sys::HB_OK {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(sys::hb_last_error()) }.to_string_lossy().into_owned();
    anyhow::bail!("hnsw_b200 error {rc}: {msg}")
}

fn as_words(s: &[Sketch]) -> *const u64 {
    s.as_ptr() as *const u64
}

/// One index resident in the HBM of one device.  `points` is the host copy of the base nodes that `Distance.point`
/// (`&'a P`, src/index/mod.rs:49-54) borrows from.
pub struct GpuIndex {
    raw: *mut sys::hb_index,
    points: Vec<Sketch>,
}
unsafe impl Send for GpuIndex {}
unsafe impl Sync for GpuIndex {} // concurrent hb_knns / hb_search calls take their own stream and workspace
impl Drop for GpuIndex {
    fn drop(&mut self) {
        unsafe { sys::hb_index_free(self.raw) }
    }
}

impl GpuIndex {
    /// `read_index` (cli/src/main.rs:224-234) straight into HBM.
    pub fn load(path: &Path, device: i32, points: Vec<Sketch>) -> anyhow::Result<Self> {
        let c = CString::new(path.to_str().ok_or_else(|| anyhow::anyhow!("non-UTF-8 path"))?)?;
        let mut raw = std::ptr::null_mut();
        check(unsafe { sys::hb_index_load_idx(c.as_ptr(), device, 0, &mut raw) })?;
        Ok(Self { raw, points })
    }

    /// `Algorithm::create` -> `HNSWBuilder::extend_parallel` (cli/src/main.rs:420-436, src/index/hnsw.rs:44-169) on the device.
    pub fn build(points: Vec<Sketch>, ef_construction: u32, connections: u32, max_connections: u32, device: i32) -> anyhow::Result<Self> {
        let p = sys::hb_build_params { ef_construction, connections, max_connections, rayon_threads: 16, serial_prefix: 0, nsw: 0 };
        let mut raw = std::ptr::null_mut();
        check(unsafe { sys::hb_build_hnsw(as_words(&points), points.len() as u64, &p, device, 1, &mut raw) })?;
        Ok(Self { raw, points })
    }

    /// `write_index` (cli/src/main.rs:236-248); needs an index built here or loaded with a host copy.
    pub fn save(&self, path: &Path) -> anyhow::Result<()> {
        let c = CString::new(path.to_str().ok_or_else(|| anyhow::anyhow!("non-UTF-8 path"))?)?;
        check(unsafe { sys::hb_index_save_idx(self.raw, c.as_ptr()) })
    }

    pub(crate) fn raw(&self) -> *const sys::hb_index {
        self.raw
    }

    fn run<'a>(&'a self, q: &[Sketch], k: usize, ef: usize) -> Vec<Vec<Distance<'a, Sketch>>> {
        let nq = q.len();
        let (mut keys, mut dists, mut found) = (vec![0u64; nq * k], vec![0u64; nq * k], vec![0u32; nq]);
        let rc = unsafe {
            sys::hb_knns(self.raw, as_words(q), nq as u64, k as u32, ef as u32, 0, keys.as_mut_ptr(), dists.as_mut_ptr(), found.as_mut_ptr(), std::ptr::null_mut())
        };
        check(rc).expect("hb_knns"); // the reference's search panics on internal errors too (src/index/nsw.rs:140)
        rows(&self.points, &keys, &dists, &found, k)
    }
}

fn rows<'a>(points: &'a [Sketch], keys: &[u64], dists: &[u64], found: &[u32], k: usize) -> Vec<Vec<Distance<'a, Sketch>>> {
    (0..found.len())
        .map(|i| {
            (0..found[i] as usize)
                .map(|j| {
                    let key = keys[i * k + j] as usize;
                    Distance::new(dists[i * k + j] as usize, key, &points[key]) // src/index/mod.rs:56-63
                })
                .collect()
        })
        .collect()
}

impl Index<Sketch> for GpuIndex {
    fn size(&self) -> usize {
        unsafe { sys::hb_index_size(self.raw) as usize } // src/index/mod.rs:22
    }

    fn search<'a>(&'a self, query: &Sketch, k: usize, ef: usize) -> Vec<Distance<'a, Sketch>> {
        self.run(std::slice::from_ref(query), k, ef).pop().unwrap() // src/index/mod.rs:23
    }

    /// ONE batched launch instead of `into_par_iter().map(search)` (src/index/mod.rs:28-42); results in input order.
    fn knns<I>(&self, queries: I, k: usize, ef: usize) -> Vec<Vec<Distance<'_, Sketch>>>
    where
        I: IntoIterator<Item = Sketch>,
    {
        let q: Vec<Sketch> = queries.into_iter().collect(); // the reference collects too (mod.rs:33)
        self.run(&q, k, ef)
    }
}

/// The same index replicated on several devices of the host; `knns` splits the batch contiguously (no collective).
pub struct MultiGpuIndex {
    raw: *mut sys::hb_multi,
    first: GpuIndex, // the replica on devices[0]; also owns `points`
}
unsafe impl Send for MultiGpuIndex {}
unsafe impl Sync for MultiGpuIndex {}
impl Drop for MultiGpuIndex {
    fn drop(&mut self) {
        unsafe { sys::hb_multi_free(self.raw) } // before `first`: the handle borrows its index
    }
}

impl MultiGpuIndex {
    pub fn replicate(first: GpuIndex, devices: &[i32]) -> anyhow::Result<Self> {
        let mut raw = std::ptr::null_mut();
        check(unsafe { sys::hb_multi_create(devices.as_ptr(), devices.len() as i32, &mut raw) })?;
        let me = Self { raw, first };
        check(unsafe { sys::hb_multi_replicate(me.raw, me.first.raw()) })?;
        Ok(me)
    }
}

impl Index<Sketch> for MultiGpuIndex {
    fn size(&self) -> usize {
        unsafe { sys::hb_multi_size(self.raw) as usize }
    }
    fn search<'a>(&'a self, query: &Sketch, k: usize, ef: usize) -> Vec<Distance<'a, Sketch>> {
        self.first.search(query, k, ef)
    }
    fn knns<I>(&self, queries: I, k: usize, ef: usize) -> Vec<Vec<Distance<'_, Sketch>>>
    where
        I: IntoIterator<Item = Sketch>,
    {
        let q: Vec<Sketch> = queries.into_iter().collect();
        let nq = q.len();
        let (mut keys, mut dists, mut found) = (vec![0u64; nq * k], vec![0u64; nq * k], vec![0u32; nq]);
        let rc = unsafe {
            sys::hb_multi_knns(self.raw, as_words(&q), nq as u64, k as u32, ef as u32, 0, keys.as_mut_ptr(), dists.as_mut_ptr(), found.as_mut_ptr())
        };
        check(rc).expect("hb_multi_knns");
        rows(&self.first.points, &keys, &dists, &found, k)
    }
}

/// The body of `GroundTruth::act` (cli/src/main.rs:715-754) without the file I/O: for every query the exact k smallest
/// (distance, key), ascending -- Bruteforce::search + MinK + the CLI's sort (src/index/bruteforce.rs:37-47,
/// src/collections/mod.rs:52-77, main.rs:738-741).  Returns 0-based keys and distances, row-major (nq, k); the CLI
/// writes `key + 1` (main.rs:745).  More than one device: rows sharded, ncclAllGather + exact merge inside the library.
pub fn ground_truth(data: &[Sketch], queries: &[Sketch], k: usize, devices: &[i32]) -> anyhow::Result<(Vec<u64>, Vec<u64>)> {
    let (mut keys, mut dists) = (vec![0u64; queries.len() * k], vec![0u64; queries.len() * k]);
    if devices.len() <= 1 {
        let dev = devices.first().copied().unwrap_or(0);
        check(unsafe {
            sys::hb_ground_truth(as_words(data), data.len() as u64, as_words(queries), queries.len() as u64, k as u32, dev, keys.as_mut_ptr(), dists.as_mut_ptr())
        })?;
    } else {
        let mut m = std::ptr::null_mut();
        check(unsafe { sys::hb_multi_create(devices.as_ptr(), devices.len() as i32, &mut m) })?;
        let rc = unsafe {
            sys::hb_multi_ground_truth(m, as_words(data), data.len() as u64, as_words(queries), queries.len() as u64, k as u32, keys.as_mut_ptr(), dists.as_mut_ptr())
        };
        unsafe { sys::hb_multi_free(m) };
        check(rc)?;
    }
    Ok((keys, dists))
}
