//! Safe wrapper over the C ABI of libb200_embed.so (include/b200_embed.h): the B200-native replacement of the tensor work in
//! `EmbeddingGenerator` (services/preprocessing_service/src/embedding_generator.rs:189-214). NOT compiled in the build image
//! of the CUDA library (no Rust toolchain there); every entry bound here is called by that repository's tests through ctypes.
use std::ffi::{c_char, c_void, CStr, CString};
use std::path::Path;

/// b200_embed_config: the config.json fields candle's bert::Config reads (embedding_generator.rs:70-71).
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct B200EmbedConfig {
    pub vocab_size: i32,
    pub hidden_size: i32,
    pub num_hidden_layers: i32,
    pub num_attention_heads: i32,
    pub intermediate_size: i32,
    pub max_position_embeddings: i32,
    pub type_vocab_size: i32,
    pub hidden_act: i32,            // 0 gelu (erf), 1 gelu_new / approximate, 2 relu
    pub layer_norm_eps: f32,
    pub max_tokens_per_launch: i32, // 0 = default
    pub operand_format: i32,        // 0 = default (fp16 tensor-core operands), 1 = bf16, 2 = fp16
}

#[repr(C)]
pub struct RawHandle { _p: [u8; 0] }
#[repr(C)]
pub struct RawIngest { _p: [u8; 0] }

extern "C" {
    fn b200_embed_create(cfg: *const B200EmbedConfig, safetensors_path: *const c_char, seed: u64, device: i32,
                         out: *mut *mut RawHandle) -> i32;
    fn b200_embed_destroy(h: *mut RawHandle);
    fn b200_embed_hidden_size(h: *const RawHandle) -> i32;
    fn b200_embed_max_positions(h: *const RawHandle) -> i32;
    fn b200_embed_last_error() -> *const c_char;
    fn b200_embed_forward(h: *mut RawHandle, ids: *const u32, mask: *const u32, type_ids: *const u32, batch: i32,
                          seq_len: i32, flags: u32, out: *mut f32) -> i32;
    fn b200_embed_forward_varlen(h: *mut RawHandle, ids: *const u32, type_ids: *const u32, cu_seqlens: *const i32,
                                 batch: i32, total_tokens: i32, max_seqlen: i32, flags: u32, out: *mut f32) -> i32;
    fn b200_embed_submit(h: *mut RawHandle, ids: *const u32, mask: *const u32, type_ids: *const u32, batch: i32,
                         seq_len: i32, flags: u32, out: *mut f32, ticket: *mut i64) -> i32;
    fn b200_embed_wait(h: *mut RawHandle, ticket: i64) -> i32;
    fn b200_embed_host_alloc(bytes: usize, out: *mut *mut c_void) -> i32;
    fn b200_embed_host_free(p: *mut c_void);
    // optional host front-end (WordPiece tokenizer.json): replaces encode_batch for BERT vocabularies
    fn b200_ingest_create(tokenizer_json_path: *const c_char, out: *mut *mut RawIngest) -> i32;
    fn b200_ingest_destroy(t: *mut RawIngest);
    fn b200_ingest_last_error() -> *const c_char;
    fn b200_ingest_encode_packed(t: *const RawIngest, text: *const c_char, begin: *const i64, end: *const i64, n: i32,
                                 max_len: i32, threads: i32, ids: *mut u32, type_ids: *mut u32, capacity_tokens: i64,
                                 cu_seqlens: *mut i32) -> i64;
}

/// Appends v / max(|v|, 1e-12). Off by default: the reference payload is not normalised (embedding_generator.rs:201-207).
pub const FLAG_NORMALIZE: u32 = 1;
/// fp32 validation mode (slow; for parity checks).
pub const FLAG_FP32: u32 = 2;

fn last_error() -> anyhow::Error {
    anyhow::anyhow!(unsafe { CStr::from_ptr(b200_embed_last_error()) }.to_string_lossy().into_owned())
}

/// One model replica on one GPU. The handle leases a stream + workspace per call behind a mutex, so it is shared like the
/// reference shares `Arc<EmbeddingGenerator>` across tokio tasks (main.rs:314, :373).
pub struct B200Embedder { h: *mut RawHandle, hidden: usize }
unsafe impl Send for B200Embedder {}
unsafe impl Sync for B200Embedder {}

impl B200Embedder {
    /// `EmbeddingGenerator::new` after the hub download (embedding_generator.rs:106-124).
    pub fn from_safetensors(cfg: &B200EmbedConfig, path: &Path, device: i32) -> anyhow::Result<Self> {
        let p = CString::new(path.to_string_lossy().as_bytes())?;
        let mut h = std::ptr::null_mut();
        if unsafe { b200_embed_create(cfg, p.as_ptr(), 0, device, &mut h) } != 0 { return Err(last_error()); }
        Ok(Self { h, hidden: unsafe { b200_embed_hidden_size(h) } as usize })
    }
    pub fn hidden_size(&self) -> usize { self.hidden }
    pub fn max_positions(&self) -> usize { unsafe { b200_embed_max_positions(self.h) as usize } }

    /// ids / mask / type_ids: `batch * seq_len` u32 each, exactly as `tokenizers` produced them (:179-187).
    pub fn forward(&self, ids: &[u32], mask: &[u32], type_ids: &[u32], batch: usize, seq_len: usize, flags: u32)
        -> anyhow::Result<Vec<Vec<f32>>> {
        anyhow::ensure!(ids.len() == batch * seq_len && mask.len() == ids.len() && type_ids.len() == ids.len());
        let mut flat = vec![0f32; batch * self.hidden];
        let rc = unsafe { b200_embed_forward(self.h, ids.as_ptr(), mask.as_ptr(), type_ids.as_ptr(), batch as i32,
                                             seq_len as i32, flags, flat.as_mut_ptr()) };
        if rc != 0 { return Err(last_error()); }
        Ok(flat.chunks(self.hidden).map(|c| c.to_vec()).collect()) // == to_vec2::<f32>() (:214)
    }

    /// Packed form: the real tokens of all sentences back to back + exclusive prefix sums (no padding crosses PCIe).
    pub fn forward_varlen(&self, ids: &[u32], type_ids: &[u32], cu_seqlens: &[i32], max_seqlen: usize, flags: u32)
        -> anyhow::Result<Vec<Vec<f32>>> {
        anyhow::ensure!(!cu_seqlens.is_empty() && ids.len() == type_ids.len());
        let batch = cu_seqlens.len() - 1;
        let mut flat = vec![0f32; batch * self.hidden];
        let rc = unsafe { b200_embed_forward_varlen(self.h, ids.as_ptr(), type_ids.as_ptr(), cu_seqlens.as_ptr(),
                                                    batch as i32, ids.len() as i32, max_seqlen as i32, flags,
                                                    flat.as_mut_ptr()) };
        if rc != 0 { return Err(last_error()); }
        Ok(flat.chunks(self.hidden).map(|c| c.to_vec()).collect())
    }

    /// Asynchronous pair for the ingest loop (main.rs:360-393): at most 4 requests in flight per handle. The buffers must
    /// stay alive (pinned, from `PinnedBuf`, for overlap) until `wait`.
    pub unsafe fn submit(&self, ids: *const u32, mask: *const u32, type_ids: *const u32, batch: usize, seq_len: usize,
                         flags: u32, out: *mut f32) -> anyhow::Result<i64> {
        let mut t = 0i64;
        if b200_embed_submit(self.h, ids, mask, type_ids, batch as i32, seq_len as i32, flags, out, &mut t) != 0 {
            return Err(last_error());
        }
        Ok(t)
    }
    pub fn wait(&self, ticket: i64) -> anyhow::Result<()> {
        if unsafe { b200_embed_wait(self.h, ticket) } != 0 { return Err(last_error()); }
        Ok(())
    }
}
impl Drop for B200Embedder { fn drop(&mut self) { unsafe { b200_embed_destroy(self.h) } } }

/// Pinned host memory (cudaHostAlloc) for the submit / wait pipeline.
pub struct PinnedBuf { p: *mut c_void, bytes: usize }
unsafe impl Send for PinnedBuf {}
impl PinnedBuf {
    pub fn new(bytes: usize) -> anyhow::Result<Self> {
        let mut p = std::ptr::null_mut();
        if unsafe { b200_embed_host_alloc(bytes, &mut p) } != 0 { return Err(last_error()); }
        Ok(Self { p, bytes })
    }
    pub fn as_mut_ptr<T>(&self) -> *mut T { self.p as *mut T }
    pub fn len_bytes(&self) -> usize { self.bytes }
}
impl Drop for PinnedBuf { fn drop(&mut self) { unsafe { b200_embed_host_free(self.p) } } }

/// Native WordPiece front-end for a BERT tokenizer.json (bit-exact with `tokenizers`, see tests/test_ingest.py of the CUDA
/// repository). Other tokenizers (Unigram / BPE) are refused at `open`: keep using the `tokenizers` crate for those.
pub struct B200Ingest { t: *mut RawIngest }
unsafe impl Send for B200Ingest {}
unsafe impl Sync for B200Ingest {}
impl B200Ingest {
    pub fn open(tokenizer_json: &Path) -> anyhow::Result<Self> {
        let p = CString::new(tokenizer_json.to_string_lossy().as_bytes())?;
        let mut t = std::ptr::null_mut();
        if unsafe { b200_ingest_create(p.as_ptr(), &mut t) } != 0 {
            anyhow::bail!(unsafe { CStr::from_ptr(b200_ingest_last_error()) }.to_string_lossy().into_owned());
        }
        Ok(Self { t })
    }
    /// `encode_batch(sentences, true)` without the padding: (ids, type_ids, cu_seqlens) for `forward_varlen`.
    pub fn encode_packed(&self, sentences: &[String], max_len: usize) -> anyhow::Result<(Vec<u32>, Vec<u32>, Vec<i32>)> {
        let mut text = Vec::new();
        let (mut begin, mut end) = (Vec::with_capacity(sentences.len()), Vec::with_capacity(sentences.len()));
        for s in sentences { begin.push(text.len() as i64); text.extend_from_slice(s.as_bytes()); end.push(text.len() as i64); }
        let mut cu = vec![0i32; sentences.len() + 1];
        let mut cap = sentences.len() * 40 + 64;
        loop {
            let (mut ids, mut tt) = (vec![0u32; cap], vec![0u32; cap]);
            let total = unsafe { b200_ingest_encode_packed(self.t, text.as_ptr() as *const c_char, begin.as_ptr(), end.as_ptr(),
                                                           sentences.len() as i32, max_len as i32, 0, ids.as_mut_ptr(),
                                                           tt.as_mut_ptr(), cap as i64, cu.as_mut_ptr()) };
            if total < 0 { anyhow::bail!(unsafe { CStr::from_ptr(b200_ingest_last_error()) }.to_string_lossy().into_owned()); }
            if (total as usize) <= cap { ids.truncate(total as usize); tt.truncate(total as usize); return Ok((ids, tt, cu)); }
            cap = total as usize;
        }
    }
}
impl Drop for B200Ingest { fn drop(&mut self) { unsafe { b200_ingest_destroy(self.t) } } }
