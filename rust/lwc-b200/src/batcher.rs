//! Request batcher: the Rust twin of `lwc::Batcher` in `llm-weighted-consensus_b200/host/lwc_host.hpp`
//! (the tested implementation).  NOT COMPILED in the build image (no Rust toolchain) -- see ../../README.md.
//!
//! Many tokio tasks call `Batcher::submit`; one OS thread per GPU owns the `Context`, groups the pending
//! requests by shape `(N, D, M)`, lays each group out `[R][N][D]` / `[R][M][N]` / `[R][M]` in pinned host
//! memory (`lwc_host_alloc`) and issues ONE `lwc_score_batch` call per group.  No tokio worker ever blocks
//! inside CUDA: `submit` copies the request (f64 -> fp32 through `lwc_stage_f64`, `Decimal` -> f32) and
//! returns a oneshot receiver that the batcher thread completes.
use crate::{response_error, Context, Scored};
use llm_weighted_consensus::{embeddings, error::ResponseError};
use lwc_b200_sys as sys;
use rust_decimal::prelude::{FromPrimitive, ToPrimitive};
use rust_decimal::Decimal;
use std::collections::HashMap;
use std::sync::mpsc;
use std::sync::Arc;
use std::time::Duration;
use tokio::sync::oneshot;

pub struct Options {
    /// requests per `lwc_score_batch` call and shape bucket
    pub max_batch: usize,
    /// how long the worker lingers for a bucket to fill
    pub max_wait: Duration,
    pub precision: i32,
}

impl Default for Options {
    fn default() -> Self {
        Options { max_batch: 4096, max_wait: Duration::from_micros(200), precision: sys::LWC_PREC_FP32 }
    }
}

struct Job {
    n: usize,
    d: usize,
    m: usize,
    e: Vec<f32>, // [N][D], empty for similarity = identity
    v: Vec<f32>, // [M][N], a NaN first element marks a voter without a vote (client.rs:410)
    w: Vec<f32>, // [M], NaN = weight None -> 0 (client.rs:412-413)
    reply: oneshot::Sender<Result<Scored, ResponseError>>,
}

pub struct Batcher {
    tx: Option<mpsc::Sender<Job>>,
    worker: Option<std::thread::JoinHandle<()>>,
    ctx: Arc<Context>,
}

/// Grow-only pinned arena, like the C++ `Pinned<T>`.
struct Pinned {
    ptr: *mut f32,
    cap: usize,
}

impl Pinned {
    fn ensure(&mut self, ctx: &Context, floats: usize) -> Result<*mut f32, ResponseError> {
        if floats > self.cap {
            unsafe {
                if !self.ptr.is_null() {
                    sys::lwc_host_free(ctx.0, self.ptr as *mut _);
                }
                let mut p = std::ptr::null_mut();
                let rc = sys::lwc_host_alloc(ctx.0, floats * 4, &mut p);
                if rc != 0 {
                    self.ptr = std::ptr::null_mut();
                    self.cap = 0;
                    return Err(response_error(rc, ctx.0));
                }
                self.ptr = p as *mut f32;
                self.cap = floats;
            }
        }
        Ok(self.ptr)
    }
}

impl Batcher {
    pub fn new(device: i32, opt: Options) -> Result<Self, ResponseError> {
        let ctx = Arc::new(Context::new(device)?);
        let (tx, rx) = mpsc::channel::<Job>();
        let worker_ctx = ctx.clone();
        let worker = std::thread::Builder::new()
            .name(format!("lwc-b200-batcher-{device}"))
            .spawn(move || run(worker_ctx, rx, opt))
            .expect("spawn batcher thread");
        Ok(Batcher { tx: Some(tx), worker: Some(worker), ctx })
    }

    pub fn context(&self) -> &Arc<Context> {
        &self.ctx
    }

    /// Thread-safe.  Everything passed in is copied before this returns.
    /// `votes[m] == None`: voter without a vote; `weights[m] == None`: 0; `embeddings == None`: `similarity = identity`.
    pub fn submit(
        &self,
        n_choices: usize,
        votes: &[Option<&[Decimal]>],
        weights: &[Option<Decimal>],
        embeddings: Option<&embeddings::response::CreateEmbeddingResponse>,
    ) -> Result<oneshot::Receiver<Result<Scored, ResponseError>>, ResponseError> {
        let bad = |msg: &str| ResponseError { code: 400, message: serde_json::Value::String(msg.to_string()) };
        let (n, m) = (n_choices, votes.len());
        if n < 1 || n > sys::LWC_MAX_CHOICES as usize {
            return Err(bad("expected 1..128 choices"));
        }
        if m > sys::LWC_MAX_VOTERS as usize || weights.len() != m {
            return Err(bad("voters and weights disagree, or more than 128 voters"));
        }
        let mut d = 0usize;
        let mut e = Vec::new();
        if let Some(resp) = embeddings {
            if resp.data.len() != n {
                return Err(bad("one embedding per choice expected"));
            }
            d = resp.data[0].embedding.len();
            e = vec![0f32; n * d];
            for row in &resp.data {
                let i = row.index as usize;
                if row.embedding.len() != d || i >= n {
                    return Err(bad("ragged embeddings"));
                }
                // Embedding.index gives the candidate row (src/embeddings/response.rs:22)
                unsafe { sys::lwc_stage_f64(self.ctx.0, row.embedding.as_ptr(), d, e[i * d..].as_mut_ptr()) };
            }
        }
        let mut v = vec![0f32; m * n];
        let mut w = vec![0f32; m];
        for (k, (vote, weight)) in votes.iter().zip(weights).enumerate() {
            w[k] = weight.and_then(|x| x.to_f32()).unwrap_or(f32::NAN);
            match vote {
                None => v[k * n..(k + 1) * n].fill(f32::NAN),
                Some(row) => {
                    if row.len() != n {
                        return Err(bad("vote length != number of choices"));
                    }
                    for (j, x) in row.iter().enumerate() {
                        v[k * n + j] = x.to_f32().unwrap_or(0.0);
                    }
                }
            }
        }
        let (reply, rx) = oneshot::channel();
        let tx = self.tx.as_ref().expect("batcher is shutting down");
        tx.send(Job { n, d, m, e, v, w, reply }).map_err(|_| ResponseError {
            code: 500,
            message: serde_json::Value::String("batcher thread is gone".into()),
        })?;
        Ok(rx)
    }
}

impl Drop for Batcher {
    fn drop(&mut self) {
        self.tx.take(); // closes the channel: the worker drains what is pending and exits
        if let Some(h) = self.worker.take() {
            let _ = h.join();
        }
    }
}

fn run(ctx: Arc<Context>, rx: mpsc::Receiver<Job>, opt: Options) {
    let mut buckets: HashMap<(usize, usize, usize), Vec<Job>> = HashMap::new();
    let mut arenas = [Pinned { ptr: std::ptr::null_mut(), cap: 0 }, Pinned { ptr: std::ptr::null_mut(), cap: 0 },
                      Pinned { ptr: std::ptr::null_mut(), cap: 0 }, Pinned { ptr: std::ptr::null_mut(), cap: 0 }];
    let mut open = true;
    while open || !buckets.is_empty() {
        // block for the first job, then linger max_wait for companions
        if buckets.is_empty() {
            match rx.recv() {
                Ok(j) => buckets.entry((j.n, j.d, j.m)).or_default().push(j),
                Err(_) => break,
            }
        }
        let deadline = std::time::Instant::now() + opt.max_wait;
        while open {
            let left = deadline.saturating_duration_since(std::time::Instant::now());
            match rx.recv_timeout(left) {
                Ok(j) => {
                    let b = buckets.entry((j.n, j.d, j.m)).or_default();
                    b.push(j);
                    if b.len() >= opt.max_batch {
                        break;
                    }
                }
                Err(mpsc::RecvTimeoutError::Timeout) => break,
                Err(mpsc::RecvTimeoutError::Disconnected) => open = false,
            }
        }
        for ((n, d, m), mut jobs) in buckets.drain() {
            while !jobs.is_empty() {
                let take = jobs.len().min(opt.max_batch);
                let batch: Vec<Job> = jobs.drain(..take).collect();
                score(&ctx, &mut arenas, &opt, n, d, m, batch);
            }
        }
    }
    for a in arenas.iter_mut() {
        if !a.ptr.is_null() {
            unsafe { sys::lwc_host_free(ctx.0, a.ptr as *mut _) };
        }
    }
}

fn reply_all(jobs: Vec<Job>, err: &ResponseError) {
    for j in jobs {
        // a dropped receiver only means the caller stopped waiting
        let _ = j.reply.send(Err(ResponseError { code: err.code, message: err.message.clone() }));
    }
}

/// One `lwc_score_batch` call for `jobs`, all of shape (n, d, m); consumes the jobs and answers each.
fn score(ctx: &Context, arenas: &mut [Pinned; 4], opt: &Options, n: usize, d: usize, m: usize, jobs: Vec<Job>) {
    let r = jobs.len();
    // arenas: 0 = E [R][N][D], 1 = V [R][M][N], 2 = w [R][M], 3 = outputs conf | cw | agree | winner | rank
    let out_floats = r * (2 * n + m.max(1)) + r + r * n;
    let mut ptrs = [std::ptr::null_mut::<f32>(); 4];
    let sizes = [(r * n * d).max(1), (r * m * n).max(1), (r * m).max(1), out_floats];
    for k in 0..4 {
        match arenas[k].ensure(ctx, sizes[k]) {
            Ok(p) => ptrs[k] = p,
            Err(e) => return reply_all(jobs, &e),
        }
    }
    let [pe, pv, pw, po] = ptrs;
    for (i, j) in jobs.iter().enumerate() {
        unsafe {
            std::ptr::copy_nonoverlapping(j.e.as_ptr(), pe.add(i * n * d), n * d);
            std::ptr::copy_nonoverlapping(j.v.as_ptr(), pv.add(i * m * n), m * n);
            std::ptr::copy_nonoverlapping(j.w.as_ptr(), pw.add(i * m), m);
        }
    }
    let desc = sys::lwc_batch_desc {
        R: r as i32, N: n as i32, D: d as i32, M: m as i32,
        similarity: if d > 0 { sys::LWC_SIM_COSINE } else { sys::LWC_SIM_IDENTITY },
        precision: opt.precision, memory: sys::LWC_MEM_HOST, w_per_request: 1,
    };
    let conf = po;
    let cw = unsafe { conf.add(r * n) };
    let agree = unsafe { cw.add(r * n) };
    let winner = unsafe { agree.add(r * m.max(1)) } as *mut i32;
    let rank = unsafe { winner.add(r) };
    let rc = unsafe {
        sys::lwc_score_batch(ctx.0, &desc, if d > 0 { pe as *const f32 } else { std::ptr::null() }, pv, pw, conf, cw, agree,
                             winner, rank, std::ptr::null_mut())
    };
    if rc != 0 {
        return reply_all(jobs, &response_error(rc, ctx.0));
    }
    let dec = |x: f32| Decimal::from_f32(x).unwrap_or(Decimal::ZERO);
    for (i, j) in jobs.into_iter().enumerate() {
        let floats = |p: *mut f32, stride: usize, len: usize| unsafe { std::slice::from_raw_parts(p.add(i * stride), len) };
        let win = unsafe { *winner.add(i) };
        let order = unsafe { std::slice::from_raw_parts(rank.add(i * n), n) };
        let scored = Scored {
            choice_weight: floats(cw, n, n).iter().map(|&x| dec(x)).collect(),
            confidence: floats(conf, n, n).iter().map(|&x| dec(x)).collect(),
            voter_confidence: floats(agree, m.max(1), m).iter().map(|&x| if x.is_nan() { None } else { Some(dec(x)) }).collect(),
            winner: if win >= 0 { Some(win as usize) } else { None },
            rank: order.iter().map(|&x| x as usize).collect(),
        };
        let _ = j.reply.send(Ok(scored));
    }
}
