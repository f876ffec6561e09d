//! assembly_helper.rs — safe wrapper with the reference's five method names over
//! shk_sys.rs.  Replaces, method for method, the wasm-bindgen `AssemblyHelper`
//! consumed at www/src/workers/Assembler.ts:15-39 for native (non-wasm) builds:
//! same constructor argument order, same JSON strings out, `Err(String)` where
//! the reference panics (Assembler.ts:101-106).  Not compiled in this image
//! (no rustc); see INTEGRATION.md.
use std::ffi::{CStr, CString};
use std::os::raw::{c_char, c_void};
use std::ptr;

use crate::shk_sys::*;

pub struct AssemblyHelper {
    h: *mut shk_asm,
}

unsafe impl Send for AssemblyHelper {} // one caller at a time, like the dedicated Web Worker

fn last_error() -> String {
    unsafe { CStr::from_ptr(shk_last_error()).to_string_lossy().into_owned() }
}
fn check(s: shk_status) -> Result<(), String> {
    if s == shk_status::SHK_OK { Ok(()) } else { Err(format!("{:?}: {}", s, last_error())) }
}

/// post_state hook (SKILLS.md:620-636): the library calls back on the calling
/// thread with the progress vocabulary of AssemblyPage.vue:451-619.
unsafe extern "C" fn trampoline(state: *const c_char, user: *mut c_void) {
    let f = &mut *(user as *mut Box<dyn FnMut(&str)>);
    f(&CStr::from_ptr(state).to_string_lossy());
}

impl AssemblyHelper {
    #[allow(clippy::too_many_arguments)]
    pub fn new(k: usize, verbose: bool, min_count: u16, min_qual: u8, chunk_size: usize, do_bloom: bool,
               do_fit: bool, no_bubble_collapse: bool, no_dead_end_removal: bool) -> Result<Self, String> {
        let p = shk_params {
            k: k as i32, verbose: verbose as i32, min_count: min_count as i32, min_qual: min_qual as i32,
            chunk_size: chunk_size as i64, do_bloom: do_bloom as i32, do_fit: do_fit as i32,
            no_bubble_collapse: no_bubble_collapse as i32, no_dead_end_removal: no_dead_end_removal as i32,
            device: -1, ..Default::default()
        };
        let mut h = ptr::null_mut();
        check(unsafe { shk_asm_new(&p, &mut h) })?;
        Ok(Self { h })
    }

    pub fn set_progress(&mut self, f: Box<dyn FnMut(&str)>) -> Result<(), String> {
        let user = Box::into_raw(Box::new(f)) as *mut c_void; // lives as long as the helper
        check(unsafe { shk_asm_set_progress_cb(self.h, Some(trampoline), user) })
    }

    /// preprocess(file1, file2|null): FASTQ, plain or gzip.
    pub fn preprocess(&mut self, file1: &str, file2: Option<&str>) -> Result<(), String> {
        let p1 = CString::new(file1).map_err(|e| e.to_string())?;
        let p2 = file2.map(CString::new).transpose().map_err(|e| e.to_string())?;
        check(unsafe { shk_asm_preprocess_files(self.h, p1.as_ptr(), p2.as_ref().map_or(ptr::null(), |c| c.as_ptr())) })
    }

    /// Streaming form for hosts that hold bytes (web_sys::File chunks, sockets).
    pub fn preprocess_push(&mut self, bytes: &[u8]) -> Result<(), String> {
        check(unsafe { shk_asm_preprocess_push(self.h, bytes.as_ptr(), bytes.len()) })
    }
    pub fn preprocess_finish(&mut self) -> Result<(), String> {
        check(unsafe { shk_asm_preprocess_finish(self.h) })
    }

    pub fn get_preprocessing_info(&self) -> Result<String, String> {
        let (mut s, mut n) = (ptr::null(), 0usize);
        check(unsafe { shk_asm_preprocessing_info(self.h, &mut s, &mut n) })?;
        Ok(unsafe { String::from_utf8_lossy(std::slice::from_raw_parts(s as *const u8, n)).into_owned() })
    }

    pub fn assemble(&mut self) -> Result<(), String> {
        check(unsafe { shk_asm_assemble(self.h) })
    }

    pub fn get_assembly(&self) -> Result<String, String> {
        let (mut s, mut n) = (ptr::null(), 0usize);
        check(unsafe { shk_asm_get_assembly(self.h, &mut s, &mut n) })?;
        Ok(unsafe { String::from_utf8_lossy(std::slice::from_raw_parts(s as *const u8, n)).into_owned() })
    }

    /// Low-level: the GPU-built graph for a host that keeps its own cleaning code.
    /// Returns (canonical k-mer lo, hi, count, validated edge byte) per node.
    pub fn kmers(&mut self) -> Result<(Vec<u64>, Vec<u64>, Vec<u16>, Vec<u8>), String> {
        let mut st = shk_stats::default();
        check(unsafe { shk_get_stats(self.h, &mut st) })?;
        let n = st.n_nodes as usize;
        let (mut lo, mut hi, mut c, mut e) = (vec![0u64; n], vec![0u64; n], vec![0u16; n], vec![0u8; n]);
        let mut got = 0usize;
        check(unsafe { shk_export_kmers(self.h, lo.as_mut_ptr(), hi.as_mut_ptr(), c.as_mut_ptr(), ptr::null_mut(),
                                        e.as_mut_ptr(), n, &mut got) })?;
        Ok((lo, hi, c, e))
    }
}

impl Drop for AssemblyHelper {
    fn drop(&mut self) {
        unsafe { shk_asm_free(self.h) } // `helper = null` in Assembler.ts:141-143
    }
}
