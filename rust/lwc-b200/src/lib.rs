//! Safe wrapper over `lwc-b200-sys` and the glue the reference crate needs for this path.
//! NOT COMPILED in the build image (no Rust toolchain) -- see ../README.md.  The tested twin of this
//! file is `llm-weighted-consensus_b200/host/lwc_host.hpp`.
use llm_weighted_consensus::{embeddings, error::ResponseError, score};
use lwc_b200_sys as sys;
use rust_decimal::prelude::{FromPrimitive, ToPrimitive};
use rust_decimal::Decimal;
use std::ffi::CStr;
use std::sync::Arc;

pub mod batcher;

/// One context per GPU.  `Send + Sync`: the library serialises calls on a context internally.
pub struct Context(*mut sys::lwc_ctx);
unsafe impl Send for Context {}
unsafe impl Sync for Context {}

fn response_error(code: i32, ctx: *const sys::lwc_ctx) -> ResponseError {
    let msg = unsafe { CStr::from_ptr(sys::lwc_last_error_message(ctx)) }.to_string_lossy().into_owned();
    // error::ResponseError { code, message } -- src/error.rs:8-13
    ResponseError { code: code as u16, message: serde_json::Value::String(msg) }
}

impl Context {
    pub fn new(device: i32) -> Result<Self, ResponseError> {
        let cfg = sys::lwc_config { device, ..Default::default() };
        let mut raw = std::ptr::null_mut();
        match unsafe { sys::lwc_ctx_create(&cfg, &mut raw) } {
            0 => Ok(Context(raw)),
            rc => Err(response_error(rc, std::ptr::null())),
        }
    }
}

impl Drop for Context {
    fn drop(&mut self) {
        unsafe { sys::lwc_ctx_destroy(self.0) }
    }
}

/// What `client.rs:422-463` writes back into the aggregate.
pub struct Scored {
    pub choice_weight: Vec<Decimal>,            // choice.weight      (client.rs:436)
    pub confidence: Vec<Decimal>,               // choice.confidence  (client.rs:437)
    pub voter_confidence: Vec<Option<Decimal>>, // client.rs:438-449
    pub winner: Option<usize>,
    pub rank: Vec<usize>,
}

/// Replaces the inline tally of `create_streaming` (client.rs:384-456) for one request.
/// `votes[m] == None`: voter without a vote (client.rs:410); `weights[m] == None`: 0 (client.rs:412-413).
/// `embeddings == None` runs `similarity = identity`, i.e. the reference arithmetic and nothing else.
pub fn score_request(
    ctx: &Context,
    n_choices: usize,
    votes: &[Option<&[Decimal]>],
    weights: &[Option<Decimal>],
    embeddings: Option<&embeddings::response::CreateEmbeddingResponse>,
) -> Result<Scored, ResponseError> {
    let (n, m) = (n_choices, votes.len());
    let mut v = vec![0f32; m * n];
    let mut w = vec![0f32; m];
    for (i, (vote, weight)) in votes.iter().zip(weights).enumerate() {
        w[i] = weight.and_then(|d| d.to_f32()).unwrap_or(f32::NAN);
        match vote {
            None => v[i * n..(i + 1) * n].fill(f32::NAN),
            Some(row) => row.iter().enumerate().for_each(|(j, d)| v[i * n + j] = d.to_f32().unwrap_or(0.0)),
        }
    }
    let mut d = 0usize;
    let mut e: Vec<f32> = Vec::new();
    if let Some(resp) = embeddings {
        d = resp.data.first().map(|x| x.embedding.len()).unwrap_or(0);
        e = vec![0f32; n * d];
        for row in &resp.data {
            // Embedding.index orders the candidates (src/embeddings/response.rs:22); Vec<f64> -> fp32
            let dst = &mut e[row.index as usize * d..(row.index as usize + 1) * d];
            unsafe { sys::lwc_stage_f64(ctx.0, row.embedding.as_ptr(), d, dst.as_mut_ptr()) };
        }
    }
    let desc = sys::lwc_batch_desc {
        R: 1, N: n as i32, D: d as i32, M: m as i32,
        similarity: if d > 0 { sys::LWC_SIM_COSINE } else { sys::LWC_SIM_IDENTITY },
        precision: sys::LWC_PREC_FP32, memory: sys::LWC_MEM_HOST, w_per_request: 1,
    };
    let (mut conf, mut cw, mut agree) = (vec![0f32; n], vec![0f32; n], vec![0f32; m.max(1)]);
    let (mut winner, mut rank) = ([0i32; 1], vec![0i32; n]);
    let rc = unsafe {
        sys::lwc_score_batch(ctx.0, &desc, if d > 0 { e.as_ptr() } else { std::ptr::null() }, v.as_ptr(), w.as_ptr(),
            conf.as_mut_ptr(), cw.as_mut_ptr(), agree.as_mut_ptr(), winner.as_mut_ptr(), rank.as_mut_ptr(),
            std::ptr::null_mut())
    };
    if rc != 0 {
        return Err(response_error(rc, ctx.0));
    }
    let dec = |x: f32| Decimal::from_f32(x).unwrap_or(Decimal::ZERO);
    Ok(Scored {
        choice_weight: cw.iter().map(|&x| dec(x)).collect(),
        confidence: conf.iter().map(|&x| dec(x)).collect(),
        voter_confidence: (0..m).map(|i| if agree[i].is_nan() { None } else { Some(dec(agree[i])) }).collect(),
        winner: if winner[0] >= 0 { Some(winner[0] as usize) } else { None },
        rank: rank.iter().map(|&x| x as usize).collect(),
    })
}

/// The same through the per-GPU batcher: the request joins whatever else is pending with its shape and the
/// caller awaits the reply without ever blocking a tokio worker inside CUDA.
pub async fn score_request_batched(
    batcher: &batcher::Batcher,
    n_choices: usize,
    votes: &[Option<&[Decimal]>],
    weights: &[Option<Decimal>],
    embeddings: Option<&embeddings::response::CreateEmbeddingResponse>,
) -> Result<Scored, ResponseError> {
    let rx = batcher.submit(n_choices, votes, weights, embeddings)?;
    rx.await.unwrap_or_else(|_| {
        Err(ResponseError { code: 500, message: serde_json::Value::String("batcher dropped the request".into()) })
    })
}

/// client.rs:386-409,459-463: `None` when some voter succeeded, else the AllVotesFailed status.
pub fn merge_error_codes(codes: &[Option<u16>]) -> Option<u16> {
    let raw: Vec<u16> = codes.iter().map(|c| c.unwrap_or(0)).collect();
    match unsafe { sys::lwc_merge_error_codes(raw.as_ptr(), raw.len() as i32) } {
        0 => None,
        c => Some(c as u16),
    }
}

impl Context {
    /// `lwc_table_load`: keep the training table (past prompt embeddings `t [rows][d]`, per-LLM quality
    /// `q [rows][m]`) resident on the GPU.
    pub fn table_load(&self, t: &[f32], q: &[f32], rows: usize, d: usize, m: usize) -> Result<(), ResponseError> {
        assert!(t.len() == rows * d && q.len() == rows * m);
        match unsafe { sys::lwc_table_load(self.0, t.as_ptr(), q.as_ptr(), rows as i64, d as i32, m as i32) } {
            0 => Ok(()),
            rc => Err(response_error(rc, self.0)),
        }
    }

    /// `lwc_table_weights` for one prompt embedding: one weight per LLM from the `top` nearest table rows.
    pub fn table_weights(&self, prompt: &[f32], top: usize, base: &[f32], min: &[f32], max: &[f32]) -> Result<Vec<f32>, ResponseError> {
        let mut w = vec![0f32; base.len()];
        let rc = unsafe {
            sys::lwc_table_weights(self.0, prompt.as_ptr(), 1, top as i32, base.as_ptr(), min.as_ptr(), max.as_ptr(),
                w.as_mut_ptr(), std::ptr::null_mut(), std::ptr::null_mut(), sys::LWC_MEM_HOST)
        };
        if rc != 0 {
            return Err(response_error(rc, self.0));
        }
        Ok(w)
    }
}

/// Fills the `FTRAININGTABLE` slot of `score::completions::Client` (client.rs:27-33), whose shipped
/// implementation is `unimplemented!()` (weight.rs:99-117).  `embed` is the user's embeddings call for the
/// prompt; the table was loaded into the context with `table_load`; `top` comes from the model's
/// `WeightTrainingTable { top, .. }` (src/score/model/mod.rs:278-306) and the limits from each LLM's
/// `base_weight / min_weight / max_weight` (src/score/llm/mod.rs:663-688).
pub struct B200TrainingTableFetcher<F> {
    pub ctx: Arc<Context>,
    pub embed: F,
}

#[async_trait::async_trait]
impl<CTX, F, Fut> score::completions::weight::Fetcher<CTX, score::completions::weight::TrainingTableData>
    for B200TrainingTableFetcher<F>
where
    CTX: Send + Sync + 'static,
    F: Fn(Arc<score::completions::request::ChatCompletionCreateParams>) -> Fut + Send + Sync,
    Fut: std::future::Future<Output = Result<embeddings::response::CreateEmbeddingResponse, ResponseError>> + Send,
{
    async fn fetch(
        &self,
        _ctx: CTX,
        request: Arc<score::completions::request::ChatCompletionCreateParams>,
        model: score::model::Model,
    ) -> Result<(Vec<Decimal>, score::completions::weight::TrainingTableData), ResponseError> {
        let embeddings_response = (self.embed)(request).await?;
        let limits: Vec<[f32; 3]> = model
            .llms
            .iter()
            .map(|llm| match llm.base.weight.weight_training_table() {
                Some(w) => [w.base_weight, w.min_weight, w.max_weight].map(|d| d.to_f32().unwrap_or(1.0)),
                None => [1.0; 3],
            })
            .collect();
        let col = |i: usize| limits.iter().map(|l| l[i]).collect::<Vec<f32>>();
        let top = match &model.weight {
            score::model::Weight::TrainingTable(t) => t.top,
            _ => 1,
        };
        let prompt: Vec<f32> = embeddings_response.data.first().map(|e| e.embedding.iter().map(|&x| x as f32).collect()).unwrap_or_default();
        // lwc_table_weights(LWC_MEM_HOST) blocks until the weights are home (a table scan of ~1 ms plus copies): it must
        // not run on a tokio worker.  The blocking pool takes it; the C++ twin (lwc_host.hpp) goes one step further and
        // lets a worker thread gather concurrent prompts into one call.
        let (ctx, base, min, max) = (self.ctx.clone(), col(0), col(1), col(2));
        let weights = tokio::task::spawn_blocking(move || ctx.table_weights(&prompt, top, &base, &min, &max))
            .await
            .map_err(|e| ResponseError { code: 500, message: serde_json::Value::String(format!("table lookup task failed: {e}")) })??;
        let weights = weights.into_iter().map(|w| Decimal::from_f32(w).unwrap_or(Decimal::ONE)).collect();
        Ok((weights, score::completions::weight::TrainingTableData { embeddings_response }))
    }
}
