//! Raw declarations of `include/lwc_b200.h`.  One item per C declaration, same order.
#![allow(non_camel_case_types, non_snake_case)]
use std::os::raw::{c_char, c_void};

pub const LWC_OK: i32 = 0;
pub const LWC_ERR_INVALID: i32 = 400;
pub const LWC_ERR_INTERNAL: i32 = 500;
pub const LWC_ERR_UNAVAILABLE: i32 = 503;
pub const LWC_MAX_CHOICES: i32 = 128;
pub const LWC_MAX_VOTERS: i32 = 128;

pub const LWC_SIM_IDENTITY: i32 = 0;
pub const LWC_SIM_COSINE: i32 = 1;
pub const LWC_PREC_FP32: i32 = 0;
pub const LWC_PREC_TF32X3: i32 = 1;
pub const LWC_PREC_TF32: i32 = 2;
pub const LWC_PREC_BF16: i32 = 3;
pub const LWC_MEM_HOST: i32 = 0;
pub const LWC_MEM_DEVICE: i32 = 1;
pub const LWC_CTX_INPUTS_RESIDENT: u64 = 1;

#[repr(C)]
pub struct lwc_ctx {
    _private: [u8; 0],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct lwc_config {
    pub device: i32,
    pub chunk_requests: i32,
    pub flags: u64,
    pub reserved: [u64; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct lwc_batch_desc {
    pub R: i32,
    pub N: i32,
    pub D: i32,
    pub M: i32,
    pub similarity: i32,
    pub precision: i32,
    pub memory: i32,
    pub w_per_request: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct lwc_pool_desc {
    pub N: i64,
    pub D: i32,
    pub M: i32,
    pub precision: i32,
    pub memory: i32,
}

pub const LWC_STATIC_SCRATCH_BYTES: u32 = 13 * 1024 * 1024;
pub const LWC_STATIC_POOL_MAX_ROWS: i64 = 16384;
pub const LWC_STATIC_POOL_MAX_ELEMS: i64 = 4194304;

extern "C" {
    pub fn lwc_ctx_create(cfg: *const lwc_config, out: *mut *mut lwc_ctx) -> i32;
    pub fn lwc_ctx_destroy(ctx: *mut lwc_ctx);
    pub fn lwc_last_error_message(ctx: *const lwc_ctx) -> *const c_char;
    pub fn lwc_version() -> *const c_char;
    pub fn lwc_host_alloc(ctx: *mut lwc_ctx, bytes: usize, out: *mut *mut c_void) -> i32;
    pub fn lwc_host_free(ctx: *mut lwc_ctx, p: *mut c_void) -> i32;
    pub fn lwc_device_alloc(ctx: *mut lwc_ctx, bytes: usize, out: *mut *mut c_void) -> i32;
    pub fn lwc_device_free(ctx: *mut lwc_ctx, p: *mut c_void) -> i32;
    pub fn lwc_memcpy(ctx: *mut lwc_ctx, dst: *mut c_void, src: *const c_void, bytes: usize, kind: i32) -> i32;
    pub fn lwc_ctx_stream(ctx: *mut lwc_ctx) -> *mut c_void;
    pub fn lwc_stream_synchronize(ctx: *mut lwc_ctx, stream: *mut c_void) -> i32;
    pub fn lwc_score_batch(
        ctx: *mut lwc_ctx,
        desc: *const lwc_batch_desc,
        E: *const f32,
        V: *const f32,
        w: *const f32,
        conf: *mut f32,
        cw: *mut f32,
        agree: *mut f32,
        winner: *mut i32,
        rank: *mut i32,
        stream: *mut c_void,
    ) -> i32;
    pub fn lwc_score_batch_typed(
        ctx: *mut lwc_ctx,
        desc: *const lwc_batch_desc,
        e_type: i32,
        E: *const c_void,
        V: *const f32,
        w: *const f32,
        conf: *mut f32,
        cw: *mut f32,
        agree: *mut f32,
        winner: *mut i32,
        rank: *mut i32,
        stream: *mut c_void,
    ) -> i32;
    pub fn lwc_stage_f64_bf16(ctx: *mut lwc_ctx, src: *const f64, n: usize, dst: *mut u16) -> i32;
    pub fn lwc_score_batches(
        ctx: *mut lwc_ctx,
        n_batches: i32,
        descs: *const lwc_batch_desc,
        E: *const *const f32,
        V: *const *const f32,
        w: *const *const f32,
        conf: *const *mut f32,
        cw: *const *mut f32,
        agree: *const *mut f32,
        winner: *const *mut i32,
        rank: *const *mut i32,
        stream: *mut c_void,
    ) -> i32;
    pub fn lwc_kernel_for(ctx: *const lwc_ctx, desc: *const lwc_batch_desc) -> *const c_char;
    pub fn lwc_launch_count(ctx: *const lwc_ctx) -> u64;
    pub fn lwc_merge_error_codes(codes: *const u16, n_voters: i32) -> i32;
    pub fn lwc_stage_f64(ctx: *mut lwc_ctx, src: *const f64, n: usize, dst: *mut f32) -> i32;
    pub fn lwc_pool_consensus(
        ctx: *mut lwc_ctx, desc: *const lwc_pool_desc, E: *const f32, V: *const f32, w: *const f32,
        conf: *mut f32, cw: *mut f32, agree: *mut f32, winner: *mut i64, stream: *mut c_void,
    ) -> i32;
    pub fn lwc_pool_consensus_static(
        ctx: *mut lwc_ctx, desc: *const lwc_pool_desc, E: *const f32, V: *const f32, w: *const f32,
        conf: *mut f32, cw: *mut f32, agree: *mut f32, winner: *mut i64, stream: *mut c_void,
    ) -> i32;
    pub fn lwc_comm_unique_id(id128: *mut c_void) -> i32;
    pub fn lwc_comm_init(ctx: *mut lwc_ctx, nranks: i32, rank: i32, id128: *const c_void) -> i32;
    pub fn lwc_comm_destroy(ctx: *mut lwc_ctx) -> i32;
    pub fn lwc_pool_p2p_export(ctx: *mut lwc_ctx, n_own: i64, D: i32, handles_out: *mut c_void) -> i32;
    pub fn lwc_pool_p2p_attach(ctx: *mut lwc_ctx, all_handles: *const c_void) -> i32;
    pub fn lwc_vote_from_logprobs(
        ctx: *mut lwc_ctx,
        n_rows: i32,
        n_top: i32,
        n_choices: i32,
        logprob: *const f32,
        leaf: *const i32,
        V: *mut f32,
        memory: i32,
        stream: *mut c_void,
    ) -> i32;
    pub fn lwc_table_load(ctx: *mut lwc_ctx, T: *const f32, Q: *const f32, rows: i64, D: i32, M: i32) -> i32;
    pub fn lwc_table_weights(
        ctx: *mut lwc_ctx,
        q: *const f32,
        n_prompts: i32,
        top: i32,
        base: *const f32,
        minw: *const f32,
        maxw: *const f32,
        weights: *mut f32,
        top_idx: *mut i32,
        top_sim: *mut f32,
        memory: i32,
    ) -> i32;
    pub fn lwc_table_merge(
        n_parts: i32,
        n_prompts: i32,
        top: i32,
        M: i32,
        row_offset: *const i64,
        idx: *const i32,
        sim: *const f32,
        q_rows: *const f32,
        base: *const f32,
        minw: *const f32,
        maxw: *const f32,
        weights: *mut f32,
        top_row: *mut i64,
        top_sim: *mut f32,
    ) -> i32;
}
