// lwc_host.hpp -- C++17 host layer above the C ABI (include/lwc_b200.h).
//
// The reference's host language is Rust; no Rust toolchain exists in the build image, so this
// header is the tested host side.  It mirrors the reference's names for THIS path so that a
// maintainer can read one against the other (paths relative to /root/reference):
//   lwc::error::ResponseError ........................ src/error.rs:8-13
//   lwc::embeddings::response::{Embedding,
//        CreateEmbeddingResponse} ..................... src/embeddings/response.rs:4-24
//   lwc::score::completions::weight::{StaticData,
//        TrainingTableData, Fetcher, StaticFetcher} ... src/score/completions/weight.rs:5-97
//   lwc::score::completions::VoterChoice ............. the fields of response::streaming::Choice the
//                                                      tally reads: delta.vote, weight, error.code
//                                                      (src/score/completions/response.rs:86-107,184-190)
//   lwc::score::completions::Tally ................... what client.rs:422-463 writes back
//   lwc::score::completions::response::streaming::
//        {Delta, Choice, ChatCompletionChunk}::push ... the aggregation of streamed chunks, path fields
//                                                      only (src/score/completions/response.rs:25-71,
//                                                      121-173,203-208; chat Delta content concat,
//                                                      src/chat/completions/response.rs:179-198)
//   lwc::score::completions::request_from_aggregate,
//        apply_tally .................................. the gather before and the write-back after the
//                                                      tally (client.rs:384-420 and 430-456)
//   lwc::Batcher ..................................... [build-defined] the packer the north star asks
//                                                      of the Rust host: many threads submit requests,
//                                                      one worker lays them out [R][N][D] in pinned
//                                                      memory per (N,D,M) bucket and calls
//                                                      lwc_score_batch; callers wait on a future, so
//                                                      no caller thread ever blocks inside CUDA.
// Error behaviour follows the reference: a failed call surfaces as ResponseError{code,message}
// (400 argument, 500 CUDA, 503 no device); nothing here falls back to the CPU.
#pragma once

#include <algorithm>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <future>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <stdexcept>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

#include "../../include/lwc_b200.h"

namespace lwc {

namespace error {
struct ResponseError : std::runtime_error {
    uint16_t code;
    std::string message;
    ResponseError(uint16_t c, std::string m) : std::runtime_error("[" + std::to_string(c) + "] " + m), code(c), message(std::move(m)) {}
};
} // namespace error

namespace embeddings::response {
struct Embedding {
    std::vector<double> embedding; // Vec<f64>, response.rs:21
    uint64_t index = 0;            // response.rs:22 -- row order of the candidates
};
struct CreateEmbeddingResponse {
    std::vector<Embedding> data;
    std::string model;
};
} // namespace embeddings::response

namespace score::completions {

namespace weight {
struct StaticData {};
struct TrainingTableData {
    embeddings::response::CreateEmbeddingResponse embeddings_response; // weight.rs:15-18
};
// trait Fetcher<CTX, T>: fetch(ctx, request, model) -> Result<(Vec<Decimal>, T), ResponseError>
// (weight.rs:66-74).  Request/model are opaque here: only the weights and the data matter to the path.
template <class CTX, class REQUEST, class MODEL, class T>
struct Fetcher {
    virtual ~Fetcher() = default;
    virtual std::pair<std::vector<double>, T> fetch(CTX ctx, const REQUEST &request, const MODEL &model) = 0;
};
} // namespace weight

struct VoterChoice {
    std::optional<std::vector<double>> vote; // Delta.vote: None -> the voter is skipped (client.rs:410)
    std::optional<double> weight;            // Choice.weight: None -> 0 (client.rs:412-413)
    std::optional<uint16_t> error_code;      // Choice.error.code (client.rs:389-408)
};

struct Tally {
    std::vector<double> weight;                          // choice.weight (client.rs:436)
    std::vector<double> confidence;                      // choice.confidence (client.rs:437)
    std::vector<std::optional<double>> voter_confidence; // client.rs:438-449; None for voters without a vote
    int32_t winner = -1;                                 // [build-defined] argmax, ties -> lowest index
    std::vector<int32_t> rank;                           // [build-defined] stable descending order
    std::optional<uint16_t> all_votes_failed;            // Error::AllVotesFailed status (client.rs:459-463)
};

struct Request {
    int32_t n_choices = 0;
    std::vector<VoterChoice> voters;
    // candidates' embeddings ordered by Embedding.index; empty -> similarity = identity (the reference)
    const embeddings::response::CreateEmbeddingResponse *embeddings = nullptr;
};

namespace response::streaming {
// Streamed chunks are merged into one aggregate before the tally (client.rs:362 `aggregate.push(&chunk)`).
// Only the fields the scoring path reads or writes are mirrored; the rule is the reference's everywhere:
// strings concatenate, every other optional keeps its FIRST value, choices merge by `index`.
struct Delta {
    std::optional<std::string> content;        // chat Delta.content: push_option_string
    std::optional<std::vector<double>> vote;   // response.rs:189,203-208: first value wins
    void push(const Delta &o) {
        if (o.content) {
            if (content) *content += *o.content;
            else content = o.content;
        }
        if (!vote) vote = o.vote;
    }
};
struct Choice {
    Delta delta;
    std::optional<std::string> finish_reason;
    uint64_t index = 0;
    std::optional<double> weight;
    std::optional<double> confidence;
    std::optional<error::ResponseError> error;
    std::optional<std::string> model;
    std::optional<size_t> model_index;
    void push(const Choice &o) { // response.rs:121-173
        delta.push(o.delta);
        if (!finish_reason) finish_reason = o.finish_reason;
        if (!weight) weight = o.weight;
        if (!confidence) confidence = o.confidence;
        if (!error) error = o.error;
        if (!model) model = o.model;
        if (!model_index) model_index = o.model_index;
    }
};
struct ChatCompletionChunk {
    std::string id;
    std::vector<Choice> choices;
    uint64_t created = 0;
    std::string model;
    void push(const ChatCompletionChunk &o) { // response.rs:25-71: choices by index, unknown index appended
        for (const Choice &oc : o.choices) {
            auto it = std::find_if(choices.begin(), choices.end(), [&](const Choice &c) { return c.index == oc.index; });
            if (it != choices.end()) it->push(oc);
            else choices.push_back(oc);
        }
    }
};
} // namespace response::streaming

// client.rs:384-420: the voters are the aggregate's choices AFTER the first `n_choices` (the request's own
// candidates come first); each contributes delta.vote, weight and error.code.
inline Request request_from_aggregate(const response::streaming::ChatCompletionChunk &aggregate, int32_t n_choices,
                                      const embeddings::response::CreateEmbeddingResponse *embeddings = nullptr) {
    Request rq;
    rq.n_choices = n_choices;
    rq.embeddings = embeddings;
    for (size_t k = (size_t)n_choices; k < aggregate.choices.size(); ++k) {
        const auto &c = aggregate.choices[k];
        VoterChoice v;
        v.vote = c.delta.vote;
        v.weight = c.weight;
        if (c.error) v.error_code = c.error->code;
        rq.voters.push_back(std::move(v));
    }
    return rq;
}

// client.rs:430-456: candidates (index < n_choices) receive weight and confidence; voters receive their
// agreement with the consensus; then delta, finish_reason and error are cleared on every choice.
inline void apply_tally(response::streaming::ChatCompletionChunk &aggregate, const Tally &tally, int32_t n_choices) {
    size_t voter = 0;
    for (size_t k = 0; k < aggregate.choices.size(); ++k) {
        auto &c = aggregate.choices[k];
        if (c.index < (uint64_t)n_choices) {
            c.weight = tally.weight[c.index];
            c.confidence = tally.confidence[c.index];
        }
        if (k >= (size_t)n_choices) { // voters are positional (skip(n)), like the gather above
            if (c.index >= (uint64_t)n_choices && voter < tally.voter_confidence.size() && tally.voter_confidence[voter])
                c.confidence = c.confidence.value_or(0.0) + *tally.voter_confidence[voter];
            ++voter;
        }
        c.delta = response::streaming::Delta{};
        c.finish_reason.reset();
        c.error.reset();
    }
}

} // namespace score::completions

// RAII over lwc_ctx.
class Context {
  public:
    explicit Context(int device = 0, int chunk_requests = 0) {
        lwc_config cfg{};
        cfg.device = device;
        cfg.chunk_requests = chunk_requests;
        const int32_t rc = lwc_ctx_create(&cfg, &ctx_);
        if (rc) throw error::ResponseError((uint16_t)rc, lwc_last_error_message(nullptr));
    }
    ~Context() { lwc_ctx_destroy(ctx_); }
    Context(const Context &) = delete;
    Context &operator=(const Context &) = delete;
    lwc_ctx *get() const { return ctx_; }
    void check(int32_t rc) const {
        if (rc) throw error::ResponseError((uint16_t)rc, lwc_last_error_message(ctx_));
    }

    // Training-table weights: what a weight::Fetcher<CTX, TrainingTableData> implementation computes
    // (weight.rs:66-74,99-117; build-defined policy, see include/lwc_b200.h).
    struct TableLimits {
        std::vector<float> base, min, max; // per LLM: base_weight / min_weight / max_weight (llm/mod.rs:663-688)
    };
    void table_load(const std::vector<float> &T, const std::vector<float> &Q, int64_t rows, int D, int M) const {
        if ((int64_t)T.size() != rows * D || (int64_t)Q.size() != rows * M) throw error::ResponseError(400, "table shape");
        check(lwc_table_load(ctx_, T.data(), Q.data(), rows, D, M));
    }
    // one weight per LLM for each of the n prompt embeddings in q [n][D]
    std::vector<float> table_weights(const std::vector<float> &q, int n, int top, const TableLimits &lim,
                                     std::vector<int32_t> *top_idx = nullptr, std::vector<float> *top_sim = nullptr) const {
        const size_t M = lim.base.size();
        std::vector<float> w((size_t)n * M);
        if (top_idx) top_idx->assign((size_t)n * top, -1);
        if (top_sim) top_sim->assign((size_t)n * top, 0.f);
        check(lwc_table_weights(ctx_, q.data(), n, top, lim.base.data(), lim.min.data(), lim.max.data(), w.data(),
                                top_idx ? top_idx->data() : nullptr, top_sim ? top_sim->data() : nullptr, LWC_MEM_HOST));
        return w;
    }

  private:
    lwc_ctx *ctx_ = nullptr;
};

// Pinned host array from lwc_host_alloc.
template <class T>
class Pinned {
  public:
    Pinned(const Context &c, size_t n) : c_(c), n_(n) {
        void *p = nullptr;
        c_.check(lwc_host_alloc(c_.get(), n * sizeof(T), &p));
        p_ = static_cast<T *>(p);
    }
    ~Pinned() { lwc_host_free(c_.get(), p_); }
    Pinned(const Pinned &) = delete;
    T *data() { return p_; }
    size_t size() const { return n_; }

  private:
    const Context &c_;
    T *p_ = nullptr;
    size_t n_;
};

// The all-votes-failed merge on the reference's own types (client.rs:386-409,459-463).
inline std::optional<uint16_t> merge_error_codes(const std::vector<score::completions::VoterChoice> &voters) {
    std::vector<uint16_t> codes;
    codes.reserve(voters.size());
    for (const auto &v : voters) codes.push_back(v.error_code.value_or(0));
    const int32_t merged = lwc_merge_error_codes(codes.data(), (int32_t)codes.size());
    if (merged == 0) return std::nullopt;
    return (uint16_t)merged;
}

class Batcher {
  public:
    struct Options {
        int max_batch = 4096;   // requests per lwc_score_batch call and bucket
        int max_wait_us = 200;  // how long the worker lingers for a bucket to fill
        int precision = LWC_PREC_FP32;
    };

    explicit Batcher(int device = 0) : Batcher(device, Options()) {}
    Batcher(int device, Options opt) : ctx_(device), opt_(opt), worker_([this] { run(); }) {}
    ~Batcher() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        worker_.join();
    }

    // Thread-safe.  The request is copied (and converted f64 -> fp32) before this returns, so the
    // caller keeps ownership of everything it passed in.
    std::future<score::completions::Tally> submit(const score::completions::Request &rq) {
        auto job = std::make_unique<Job>();
        const int N = rq.n_choices, M = (int)rq.voters.size();
        if (N < 1 || N > LWC_MAX_CHOICES) throw error::ResponseError(400, "expected 1.." + std::to_string(LWC_MAX_CHOICES) + " choices");
        if (M > LWC_MAX_VOTERS) throw error::ResponseError(400, "more than " + std::to_string(LWC_MAX_VOTERS) + " voters");
        int D = 0;
        if (rq.embeddings) {
            if ((int)rq.embeddings->data.size() != N) throw error::ResponseError(400, "one embedding per choice expected");
            D = (int)rq.embeddings->data[0].embedding.size();
            job->E.resize((size_t)N * D);
            for (const auto &e : rq.embeddings->data) {
                if ((int)e.embedding.size() != D || e.index >= (uint64_t)N) throw error::ResponseError(400, "ragged embeddings");
                // Embedding.index gives the candidate row (response.rs:22); f64 -> fp32 (lwc_stage_f64)
                lwc_stage_f64(ctx_.get(), e.embedding.data(), (size_t)D, job->E.data() + (size_t)e.index * D);
            }
        }
        job->N = N, job->D = D, job->M = M;
        job->V.assign((size_t)M * N, 0.f);
        job->w.assign((size_t)M, 0.f);
        const float nan = std::numeric_limits<float>::quiet_NaN();
        for (int m = 0; m < M; ++m) {
            const auto &vc = rq.voters[m];
            job->w[m] = vc.weight ? (float)*vc.weight : nan;
            if (!vc.vote) {
                std::fill_n(job->V.begin() + (size_t)m * N, N, nan);
            } else {
                if ((int)vc.vote->size() != N) throw error::ResponseError(400, "vote length != number of choices");
                for (int i = 0; i < N; ++i) job->V[(size_t)m * N + i] = (float)(*vc.vote)[i];
            }
        }
        job->failed = merge_error_codes(rq.voters);
        auto fut = job->promise.get_future();
        {
            std::lock_guard<std::mutex> lk(mu_);
            queue_[Key{N, D, M}].push_back(std::move(job));
            ++pending_;
        }
        cv_.notify_one();
        return fut;
    }

    const Context &context() const { return ctx_; }
    uint64_t batches() const { return batches_; }

  private:
    struct Job {
        int N = 0, D = 0, M = 0;
        std::vector<float> E, V, w;
        std::optional<uint16_t> failed;
        std::promise<score::completions::Tally> promise;
    };
    using Key = std::tuple<int, int, int>;

    void run() {
        for (;;) {
            std::vector<std::unique_ptr<Job>> batch;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [this] { return stop_ || pending_ > 0; });
                if (stop_ && pending_ == 0) return;
                // linger briefly so that concurrent submitters land in the same launch
                if (!stop_) cv_.wait_for(lk, std::chrono::microseconds(opt_.max_wait_us), [this] { return stop_ || pending_ >= opt_.max_batch; });
                // take the fullest bucket
                auto best = queue_.end();
                for (auto it = queue_.begin(); it != queue_.end(); ++it)
                    if (!it->second.empty() && (best == queue_.end() || it->second.size() > best->second.size())) best = it;
                if (best == queue_.end()) continue;
                while (!best->second.empty() && (int)batch.size() < opt_.max_batch) {
                    batch.push_back(std::move(best->second.front()));
                    best->second.pop_front();
                    --pending_;
                }
            }
            score(batch);
        }
    }

    void score(std::vector<std::unique_ptr<Job>> &batch) {
        const int R = (int)batch.size(), N = batch[0]->N, D = batch[0]->D, M = batch[0]->M;
        try {
            // pinned staging is kept across batches and only ever grows: cudaHostAlloc costs milliseconds
            Arena E(arena_[0], ctx_, (size_t)R * N * std::max(D, 1)), V(arena_[1], ctx_, (size_t)R * std::max(M, 1) * N),
                w(arena_[2], ctx_, (size_t)R * std::max(M, 1));
            Arena conf(arena_[3], ctx_, (size_t)R * N), cw(arena_[4], ctx_, (size_t)R * N),
                agree(arena_[5], ctx_, (size_t)R * std::max(M, 1));
            ArenaI winner(arena_[6], ctx_, (size_t)R), rank(arena_[7], ctx_, (size_t)R * N);
            for (int r = 0; r < R; ++r) {
                if (D) std::memcpy(E.data() + (size_t)r * N * D, batch[r]->E.data(), sizeof(float) * N * D);
                if (M) std::memcpy(V.data() + (size_t)r * M * N, batch[r]->V.data(), sizeof(float) * M * N);
                if (M) std::memcpy(w.data() + (size_t)r * M, batch[r]->w.data(), sizeof(float) * M);
            }
            lwc_batch_desc d{};
            d.R = R, d.N = N, d.D = D, d.M = M;
            d.similarity = D ? LWC_SIM_COSINE : LWC_SIM_IDENTITY;
            d.precision = opt_.precision;
            d.memory = LWC_MEM_HOST;
            d.w_per_request = 1;
            ctx_.check(lwc_score_batch(ctx_.get(), &d, D ? E.data() : nullptr, V.data(), w.data(), conf.data(), cw.data(),
                                       agree.data(), winner.data(), rank.data(), nullptr));
            ++batches_;
            for (int r = 0; r < R; ++r) {
                score::completions::Tally t;
                t.weight.assign(cw.data() + (size_t)r * N, cw.data() + (size_t)(r + 1) * N);
                t.confidence.assign(conf.data() + (size_t)r * N, conf.data() + (size_t)(r + 1) * N);
                t.rank.assign(rank.data() + (size_t)r * N, rank.data() + (size_t)(r + 1) * N);
                t.winner = winner.data()[r];
                t.voter_confidence.resize(M);
                for (int m = 0; m < M; ++m) {
                    const float a = agree.data()[(size_t)r * M + m];
                    if (!std::isnan(a)) t.voter_confidence[m] = a;
                }
                t.all_votes_failed = batch[r]->failed;
                batch[r]->promise.set_value(std::move(t));
            }
        } catch (...) {
            for (auto &j : batch) {
                try {
                    j->promise.set_exception(std::current_exception());
                } catch (...) {
                }
            }
        }
    }

    // view of a grow-only pinned buffer (one per role; touched by the worker thread only)
    template <class T>
    struct ArenaT {
        ArenaT(std::unique_ptr<Pinned<T>> &slot, const Context &c, size_t n) {
            if (!slot || slot->size() < n) slot = std::make_unique<Pinned<T>>(c, n + n / 4 + 64);
            p = slot->data();
        }
        T *data() { return p; }
        T *p;
    };
    struct Arena : ArenaT<float> {
        Arena(std::unique_ptr<Pinned<float>> &s, const Context &c, size_t n) : ArenaT<float>(s, c, n) {}
    };
    struct ArenaI {
        ArenaI(std::unique_ptr<Pinned<float>> &s, const Context &c, size_t n) : a(s, c, n) {}
        int32_t *data() { return reinterpret_cast<int32_t *>(a.data()); }
        ArenaT<float> a;
    };

    Context ctx_;
    Options opt_;
    std::unique_ptr<Pinned<float>> arena_[8];
    std::mutex mu_;
    std::condition_variable cv_;
    std::map<Key, std::deque<std::unique_ptr<Job>>> queue_;
    int pending_ = 0;
    bool stop_ = false;
    uint64_t batches_ = 0;
    std::thread worker_;
};

} // namespace lwc
