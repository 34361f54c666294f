// lwc_host.hpp -- C++17 host layer above the C ABI (include/lwc_b200.h).
//
// The reference's host language is Rust; no Rust toolchain exists in the build image, so this
// header is the tested host side.  It mirrors the reference's names for THIS path so that a
// maintainer can read one against the other (paths relative to /root/reference):
//   lwc::error::ResponseError ........................ src/error.rs:8-13
//   lwc::embeddings::response::{Embedding,
//        CreateEmbeddingResponse} ..................... src/embeddings/response.rs:4-24
//   lwc::score::{WeightType, llm::Weight, model::{Llm,
//        Weight, Model}} .............................. the weight configuration the fetchers read: src/score/mod.rs:7-13,
//                                                      src/score/llm/mod.rs:590-688, src/score/model/mod.rs:278-306
//   lwc::score::completions::weight::{StaticData,
//        TrainingTableData, Data, Fetcher, Fetchers,
//        StaticFetcher} ............................... src/score/completions/weight.rs:5-97
//   lwc::score::completions::weight::
//        B200TrainingTableFetcher ..................... the slot weight.rs:99-117 leaves unimplemented!(): prompt
//                                                      embedding -> lwc_table_weights, batched by a worker thread
//   lwc::chat::completions::response::unary::Message,
//   lwc::multichat::completions::response::unary::
//        {Choice, ChatCompletion} ..................... the candidate source types, src/multichat/completions/
//                                                      response.rs:150-197 (text fields only)
//   lwc::score::completions::request::{Choice,
//        InternalChoice, convert_choices_to_internal_choices,
//        choice_texts, choices_from_multichat} ........ which text -- hence which row of E -- each candidate is:
//                                                      request.rs:68-110, client.rs:144-170,1078-1160,1222-1289
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
#include <atomic>
#include <chrono>
#include <deque>
#include <functional>
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

namespace chat::completions::response::unary {
// chat Message, text fields only (src/chat/completions/response.rs; tool calls are rendered by the reference
// with serde_json and are not part of this mirror)
struct Message {
    std::optional<std::string> content, refusal, reasoning;
};
} // namespace chat::completions::response::unary

namespace multichat::completions::response::unary {
// src/multichat/completions/response.rs:184-197 / 150-158: one generation per LLM of a multichat model
struct Choice {
    chat::completions::response::unary::Message message;
    uint64_t index = 0;
    std::optional<error::ResponseError> error;
    std::optional<std::string> model;
    std::optional<size_t> model_index;
};
struct ChatCompletion {
    std::string id;
    std::vector<Choice> choices;
    uint64_t created = 0;
    std::string model;
};
} // namespace multichat::completions::response::unary

namespace score {
enum class WeightType { Static, TrainingTable }; // src/score/mod.rs:7-13
namespace llm {
// src/score/llm/mod.rs:590-688: a voter's weight is one number, or base / min / max for the training-table mode
struct Weight {
    WeightType type = WeightType::Static;
    double weight = 1.0;                                       // WeightStatic.weight (default dec!(1.0), llm/mod.rs:602)
    double base_weight = 1.0, min_weight = 1.0, max_weight = 1.0; // WeightTrainingTable
    // llm/mod.rs:650-660, 671-687
    std::optional<std::string> validate() const {
        if (type == WeightType::Static) {
            if (!(weight > 0)) return "`weight` must be a normal positive number";
        } else if (base_weight < min_weight || base_weight > max_weight || min_weight > max_weight || !(base_weight > 0) ||
                   !(min_weight > 0) || !(max_weight > 0)) {
            return "LLM must have normal positive base, min, and max weights for training table weights mode";
        }
        return std::nullopt;
    }
};
} // namespace llm
namespace model {
struct Llm {
    std::string id;
    llm::Weight weight; // llm.base.weight
    size_t index = 0;   // position after the reference's sort-by-id (model/mod.rs:88-94,134-137): weights[llm.index]
};
// model/mod.rs:278-306: Static, or TrainingTable { embeddings { model, max_tokens }, top }
struct Weight {
    WeightType type = WeightType::Static;
    std::string embeddings_model;
    size_t embeddings_max_tokens = 0;
    size_t top = 0;
};
struct Model {
    std::string id;
    std::vector<Llm> llms;
    Weight weight;
};
} // namespace model
} // namespace score

namespace score::completions {

namespace request {
// request.rs:68-91: a candidate is a literal text, a message, or a reference to a choice of an archived completion
struct Choice {
    enum class Kind { Text, ChatCompletion, ScoreCompletion, MultichatCompletion, ChatCompletionChoiceMessage } kind = Kind::Text;
    std::string text;          // Text
    std::string id;            // *Completion: id of the archived completion
    uint64_t choice_index = 0; // *Completion: which of its choices (serde default 0)
    chat::completions::response::unary::Message message; // ChatCompletionChoiceMessage
    static Choice Text(std::string t) {
        Choice c;
        c.text = std::move(t);
        return c;
    }
    static Choice MultichatCompletion(std::string completion_id, uint64_t index) {
        Choice c;
        c.kind = Kind::MultichatCompletion;
        c.id = std::move(completion_id);
        c.choice_index = index;
        return c;
    }
};
// request.rs:93-110, the variants this path can resolve without the chat / score archives
struct InternalChoice {
    enum class Kind { Text, ChatCompletionChoiceMessage, MultichatCompletionChoice } kind = Kind::Text;
    std::string text;
    chat::completions::response::unary::Message message;
    multichat::completions::response::unary::Choice multichat_choice;
};
// the multichat part of completions_archive::Completion (fetched by id, client.rs:990-1001)
using MultichatArchive = std::map<std::string, multichat::completions::response::unary::ChatCompletion>;

// client.rs:1222-1289 without tool calls: reasoning, content, refusal, separated by blank lines
inline std::string convert_completion_message_to_text(const chat::completions::response::unary::Message &m) {
    std::string text;
    for (const auto *part : {&m.reasoning, &m.content, &m.refusal})
        if (*part) {
            if (!text.empty()) text += "\n\n";
            text += **part;
        }
    return text;
}

// client.rs:1078-1160.  The i-th element of the result is the i-th element of `choices`: candidate order is the
// request's order, and it is the ROW ORDER of E -- the embedding of candidate i must be sent with
// Embedding.index = i (embeddings/response.rs:22).  A reference to a choice the archived completion does not
// hold is Error::InvalidCompletionChoiceIndex -> 400 (score/completions/error.rs:22,43).
inline std::vector<InternalChoice> convert_choices_to_internal_choices(const MultichatArchive &completions,
                                                                       const std::vector<Choice> &choices) {
    std::vector<InternalChoice> out;
    out.reserve(choices.size());
    for (const Choice &c : choices) {
        InternalChoice ic;
        switch (c.kind) {
            case Choice::Kind::Text:
                ic.kind = InternalChoice::Kind::Text;
                ic.text = c.text;
                break;
            case Choice::Kind::ChatCompletionChoiceMessage:
                ic.kind = InternalChoice::Kind::ChatCompletionChoiceMessage;
                ic.message = c.message;
                break;
            case Choice::Kind::MultichatCompletion: {
                const auto it = completions.find(c.id);
                if (it == completions.end()) throw error::ResponseError(404, "completion " + c.id + " is not in the archive");
                const auto &cs = it->second.choices;
                const auto ch = std::find_if(cs.begin(), cs.end(), [&](const auto &x) { return x.index == c.choice_index; });
                if (ch == cs.end())
                    throw error::ResponseError(400, "invalid choice_index for completion " + c.id + ": " + std::to_string(c.choice_index));
                ic.kind = InternalChoice::Kind::MultichatCompletionChoice;
                ic.multichat_choice = *ch;
                break;
            }
            default:
                throw error::ResponseError(400, "chat / score completion archives are outside this path's mirror");
        }
        out.push_back(std::move(ic));
    }
    return out;
}

// client.rs:149-170: the text each candidate is scored (and embedded) as, in row order
inline std::vector<std::string> choice_texts(const std::vector<InternalChoice> &internal) {
    std::vector<std::string> texts;
    texts.reserve(internal.size());
    for (const auto &ic : internal)
        texts.push_back(ic.kind == InternalChoice::Kind::Text ? ic.text
                        : ic.kind == InternalChoice::Kind::ChatCompletionChoiceMessage
                            ? convert_completion_message_to_text(ic.message)
                            : convert_completion_message_to_text(ic.multichat_choice.message));
    return texts;
}

// [build-defined] "batching requests from src/multichat": every generation of one multichat completion becomes a
// candidate of one score request, in the completion's own choice order (failed generations are left out).
inline std::vector<Choice> choices_from_multichat(const multichat::completions::response::unary::ChatCompletion &mc) {
    std::vector<Choice> out;
    for (const auto &c : mc.choices)
        if (!c.error) out.push_back(Choice::MultichatCompletion(mc.id, c.index));
    return out;
}
} // namespace request

namespace weight {
struct StaticData {};
struct TrainingTableData {
    embeddings::response::CreateEmbeddingResponse embeddings_response; // weight.rs:15-18
};
// enum Data { Static(StaticData), TrainingTable(TrainingTableData) }, weight.rs:5-10
struct Data {
    WeightType type = WeightType::Static;
    TrainingTableData training_table;
};
// trait Fetcher<CTX, T>: async fn fetch(ctx, request, model) -> Result<(Vec<Decimal>, T), ResponseError>
// (weight.rs:66-74).  `async` becomes a future: fetch must not block its caller inside CUDA, errors
// travel in the future as ResponseError.  The request is opaque here: only what an embedder needs of it matters.
template <class CTX, class REQUEST, class T>
struct Fetcher {
    virtual ~Fetcher() = default;
    virtual std::future<std::pair<std::vector<double>, T>> fetch(CTX ctx, std::shared_ptr<const REQUEST> request,
                                                                 const model::Model &model) = 0;
};
// weight.rs:76-97: the configured per-LLM weights, in llm order
template <class CTX, class REQUEST>
struct StaticFetcher : Fetcher<CTX, REQUEST, StaticData> {
    std::future<std::pair<std::vector<double>, StaticData>> fetch(CTX, std::shared_ptr<const REQUEST>,
                                                                  const model::Model &model) override {
        std::promise<std::pair<std::vector<double>, StaticData>> p;
        std::vector<double> w;
        w.reserve(model.llms.size());
        for (const auto &llm : model.llms) w.push_back(llm.weight.weight); // weight_static().unwrap().weight
        p.set_value({std::move(w), StaticData{}});
        return p.get_future();
    }
};
// weight.rs:21-63: dispatch on model.weight.type()
template <class CTX, class REQUEST>
struct Fetchers {
    std::shared_ptr<Fetcher<CTX, REQUEST, StaticData>> static_;
    std::shared_ptr<Fetcher<CTX, REQUEST, TrainingTableData>> training_table;
    std::pair<std::vector<double>, Data> fetch(CTX ctx, std::shared_ptr<const REQUEST> request, const model::Model &model) const {
        Data d;
        d.type = model.weight.type;
        if (model.weight.type == WeightType::Static) {
            auto r = static_->fetch(ctx, std::move(request), model).get();
            return {std::move(r.first), std::move(d)};
        }
        auto r = training_table->fetch(ctx, std::move(request), model).get();
        d.training_table = std::move(r.second);
        return {std::move(r.first), std::move(d)};
    }
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

// The producer side of lwc_vote_from_logprobs: the part of get_vote that stays on the host.  The reference walks the
// top-logprob alternatives of the token that carries the final prefix character (client.rs:1766-1784) and, for each
// alternative that (a) is long enough to have a character at that byte offset, (b) has a logprob, (c) has a prefix
// letter there, (d) whose letter is a sibling leaf of the lowest branch, adds exp(logprob) to vote[leaf].  Conditions
// (a)-(d) are string work and stay here; what is handed to the device is one (logprob, leaf) pair per alternative,
// leaf = -1 when any condition fails.  `leaf_of` is pfx_tree.get(&pfx).unwrap_leaf() for the voter's tree.
namespace chat::completions::response {
struct TopLogprob {
    std::string token;
    std::optional<double> logprob; // TopLogprob.logprob: Option<Decimal>
};
} // namespace chat::completions::response

namespace score::completions {
// Writes n_top pairs at lp[0..n_top) / leaf[0..n_top) (row `voter` of the [rows][n_top] arrays the C entry takes);
// alternatives beyond n_top are ignored like the reference ignores what the upstream did not return (top_logprobs
// <= 20, llm/mod.rs:455-467), missing ones are padded with leaf = -1.
inline void fill_vote_pairs(const std::vector<chat::completions::response::TopLogprob> &top_logprobs, size_t key_logprob_index,
                            const std::function<std::optional<int32_t>(char)> &leaf_of, int n_top, float *lp, int32_t *leaf) {
    const float nan = std::numeric_limits<float>::quiet_NaN();
    for (int k = 0; k < n_top; ++k) {
        lp[k] = nan;
        leaf[k] = -1;
        if (k >= (int)top_logprobs.size()) continue;
        const auto &alt = top_logprobs[(size_t)k];
        if (key_logprob_index >= alt.token.size() || !alt.logprob) continue;          // client.rs:1772-1773
        const unsigned char c = (unsigned char)alt.token[key_logprob_index];
        if (c >= 0x80) continue; // char_indices().find(|(i, _)| *i == index): a continuation byte is not a char start,
                                 // and no multi-byte char is a prefix letter (A..T, client.rs:1343-1364)
        const auto l = leaf_of((char)c);                                              // client.rs:1777-1778
        if (!l) continue;
        lp[k] = (float)*alt.logprob;
        leaf[k] = *l;
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

// weight::Fetcher<CTX, TrainingTableData> for WeightType::TrainingTable -- the slot the reference ships as
// unimplemented!() (weight.rs:99-117).  `embed` is the caller's embeddings client: it turns the request's prompt
// into a CreateEmbeddingResponse with ONE row (that response is also what TrainingTableData carries back,
// weight.rs:15-18).  fetch() runs `embed` on the calling thread, hands the row to a worker thread and returns a
// future: the worker gathers the prompts that arrived meanwhile into one lwc_table_weights call (the score pass
// streams the whole resident table once per call, eight prompts per pass), so no caller ever blocks inside CUDA.
// The table itself is loaded into `ctx` beforehand (Context::table_load); `top` and the per-LLM base / min / max
// weights come from the model (model/mod.rs:278-306, llm/mod.rs:663-688).
namespace score::completions::weight {
template <class CTX, class REQUEST>
class B200TrainingTableFetcher : public Fetcher<CTX, REQUEST, TrainingTableData> {
  public:
    using Embed = std::function<embeddings::response::CreateEmbeddingResponse(CTX, const REQUEST &, const model::Model &)>;
    using Result = std::pair<std::vector<double>, TrainingTableData>;

    B200TrainingTableFetcher(const Context &ctx, int table_dim, Embed embed, int max_prompts_per_call = 64)
        : ctx_(ctx), D_(table_dim), embed_(std::move(embed)), max_call_(max_prompts_per_call), worker_([this] { run(); }) {}
    ~B200TrainingTableFetcher() override {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        worker_.join();
    }

    std::future<Result> fetch(CTX ctx, std::shared_ptr<const REQUEST> request, const model::Model &model) override {
        auto job = std::make_unique<Job>();
        auto fut = job->promise.get_future();
        try {
            if (model.weight.type != WeightType::TrainingTable) throw error::ResponseError(400, "expected weight of type `training_table`");
            if (model.weight.top < 1 || model.weight.top > 64) throw error::ResponseError(400, "training table weight `top` must be 1..64 here");
            job->top = (int)model.weight.top;
            for (const auto &llm : model.llms) {
                if (auto bad = llm.weight.validate()) throw error::ResponseError(400, *bad);
                job->base.push_back((float)llm.weight.base_weight);
                job->minw.push_back((float)llm.weight.min_weight);
                job->maxw.push_back((float)llm.weight.max_weight);
            }
            job->data.embeddings_response = embed_(ctx, *request, model);
            const auto &rows = job->data.embeddings_response.data;
            if (rows.size() != 1 || (int)rows[0].embedding.size() != D_)
                throw error::ResponseError(500, "the embeddings client must return one row of the table's dimension");
            job->q.resize((size_t)D_);
            lwc_stage_f64(ctx_.get(), rows[0].embedding.data(), (size_t)D_, job->q.data());
        } catch (...) {
            job->promise.set_exception(std::current_exception());
            return fut;
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            queue_.push_back(std::move(job));
        }
        cv_.notify_one();
        return fut;
    }
    uint64_t calls() const { return calls_; }

  private:
    struct Job {
        int top = 0;
        std::vector<float> base, minw, maxw, q;
        TrainingTableData data;
        std::promise<Result> promise;
        bool same_policy(const Job &o) const { return top == o.top && base == o.base && minw == o.minw && maxw == o.maxw; }
    };
    void run() {
        for (;;) {
            std::vector<std::unique_ptr<Job>> batch;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [this] { return stop_ || !queue_.empty(); });
                if (queue_.empty()) return;
                // one call per policy (model): take the head and everything queued behind it with the same limits
                batch.push_back(std::move(queue_.front()));
                queue_.pop_front();
                for (auto it = queue_.begin(); it != queue_.end() && (int)batch.size() < max_call_;)
                    if ((*it)->same_policy(*batch[0])) {
                        batch.push_back(std::move(*it));
                        it = queue_.erase(it);
                    } else {
                        ++it;
                    }
            }
            const int n = (int)batch.size(), M = (int)batch[0]->base.size();
            try {
                std::vector<float> q((size_t)n * D_), w((size_t)n * M);
                for (int i = 0; i < n; ++i) std::memcpy(q.data() + (size_t)i * D_, batch[i]->q.data(), sizeof(float) * D_);
                ctx_.check(lwc_table_weights(ctx_.get(), q.data(), n, batch[0]->top, batch[0]->base.data(), batch[0]->minw.data(),
                                             batch[0]->maxw.data(), w.data(), nullptr, nullptr, LWC_MEM_HOST));
                ++calls_;
                for (int i = 0; i < n; ++i)
                    batch[i]->promise.set_value({std::vector<double>(w.begin() + (size_t)i * M, w.begin() + (size_t)(i + 1) * M),
                                                 std::move(batch[i]->data)});
            } catch (...) {
                for (auto &j : batch) j->promise.set_exception(std::current_exception());
            }
        }
    }
    const Context &ctx_;
    int D_;
    Embed embed_;
    int max_call_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<std::unique_ptr<Job>> queue_;
    bool stop_ = false;
    std::atomic<uint64_t> calls_{0};
    std::thread worker_;
};
} // namespace score::completions::weight

// [build-defined] The packer the north star asks of the host: many threads submit requests; each request is
// converted (f64 -> fp32, candidates placed by Embedding.index) by its OWN submitting thread STRAIGHT INTO its
// slot of a pinned, device-contiguous [R][N][D] arena -- one pass over every embedding, no intermediate copy --
// and a worker thread hands full (or lingering) arenas to lwc_score_batch.  Every (N, D, M) shape has its own
// bucket with two arenas, so the submitters fill one while the GPU works through the other.  Callers wait on a
// future and never block inside CUDA; submit() blocks only when both arenas of its bucket are in use
// (back-pressure).
class Batcher {
  public:
    struct Options {
        int max_batch = 4096;            // requests per lwc_score_batch call and bucket
        int max_wait_us = 200;           // how long a partly filled arena lingers for more requests
        int precision = LWC_PREC_FP32;
        size_t max_arena_bytes = 512u << 20; // pinned bytes of embeddings per arena (caps max_batch for large shapes)
        // element type the embeddings cross PCIe as (include/lwc_b200.h, lwc_score_batch_typed): LWC_ELEM_F32 converts
        // f64 -> fp32 on the submitting threads; LWC_ELEM_F64 copies the doubles and lets the device round them;
        // LWC_ELEM_BF16 (opt-in, 1e-2 tolerance) converts f64 -> bf16 on the submitting threads
        int transport = LWC_ELEM_F32;
    };

    explicit Batcher(int device = 0) : Batcher(device, Options()) {}
    Batcher(int device, Options opt) : ctx_(device), opt_(opt), worker_([this] { run(); }) {}
    ~Batcher() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_work_.notify_all();
        worker_.join();
    }

    // Thread-safe.  The request is copied (and converted f64 -> fp32) before this returns, so the
    // caller keeps ownership of everything it passed in.
    std::future<score::completions::Tally> submit(const score::completions::Request &rq) {
        const int N = rq.n_choices, M = (int)rq.voters.size();
        if (N < 1 || N > LWC_MAX_CHOICES) throw error::ResponseError(400, "expected 1.." + std::to_string(LWC_MAX_CHOICES) + " choices");
        if (M > LWC_MAX_VOTERS) throw error::ResponseError(400, "more than " + std::to_string(LWC_MAX_VOTERS) + " voters");
        int D = 0;
        if (rq.embeddings) {
            if ((int)rq.embeddings->data.size() != N) throw error::ResponseError(400, "one embedding per choice expected");
            D = (int)rq.embeddings->data[0].embedding.size();
            std::vector<bool> seen((size_t)N, false);
            for (const auto &e : rq.embeddings->data) {
                if ((int)e.embedding.size() != D || e.index >= (uint64_t)N || seen[e.index]) throw error::ResponseError(400, "ragged embeddings");
                seen[e.index] = true;
            }
        }
        for (const auto &vc : rq.voters)
            if (vc.vote && (int)vc.vote->size() != N) throw error::ResponseError(400, "vote length != number of choices");

        // ---- reserve a slot of the bucket's open arena
        Arena *a = nullptr;
        int slot = 0;
        std::future<score::completions::Tally> fut;
        {
            std::unique_lock<std::mutex> lk(mu_);
            Bucket &b = bucket_for(N, D, M);
            for (;;) {
                Arena &cur = b.arena[b.open];
                if (cur.state == Arena::Filling && cur.reserved < b.cap) break;
                Arena &other = b.arena[b.open ^ 1];
                if (other.state == Arena::Filling && other.reserved < b.cap) {
                    b.open ^= 1;
                    break;
                }
                cv_space_.wait(lk); // both arenas sealed or being scored
            }
            a = &b.arena[b.open];
            slot = a->reserved++;
            if (slot == 0) a->first = std::chrono::steady_clock::now();
            a->promises.emplace_back();
            fut = a->promises.back().get_future();
            a->failed.push_back(merge_error_codes(rq.voters));
            if (a->reserved == b.cap) a->state = Arena::Sealed;
        }
        cv_work_.notify_one();

        // ---- fill the slot outside the lock: the only pass over this request's embeddings
        if (D)
            for (const auto &e : rq.embeddings->data) { // Embedding.index gives the candidate row (response.rs:22)
                const size_t at = ((size_t)slot * N + e.index) * D;
                if (opt_.transport == LWC_ELEM_F64)
                    std::memcpy(reinterpret_cast<double *>(a->E->data()) + at, e.embedding.data(), sizeof(double) * D);
                else if (opt_.transport == LWC_ELEM_BF16)
                    lwc_stage_f64_bf16(ctx_.get(), e.embedding.data(), (size_t)D, reinterpret_cast<uint16_t *>(a->E->data()) + at);
                else
                    lwc_stage_f64(ctx_.get(), e.embedding.data(), (size_t)D, reinterpret_cast<float *>(a->E->data()) + at);
            }
        const float nan = std::numeric_limits<float>::quiet_NaN();
        float *V = a->V->data() + (size_t)slot * M * N, *w = a->w->data() + (size_t)slot * M;
        for (int m = 0; m < M; ++m) {
            const auto &vc = rq.voters[m];
            w[m] = vc.weight ? (float)*vc.weight : nan;
            if (!vc.vote)
                std::fill_n(V + (size_t)m * N, N, nan);
            else
                for (int i = 0; i < N; ++i) V[(size_t)m * N + i] = (float)(*vc.vote)[i];
        }
        a->ready.fetch_add(1, std::memory_order_release);
        if (a->reserved_at_seal.load(std::memory_order_acquire) >= 0) { // the worker already took this arena and may be
            { std::lock_guard<std::mutex> lk(mu_); }                    // waiting for this slot: no lost wake-up
            cv_work_.notify_one();
        }
        return fut;
    }

    const Context &context() const { return ctx_; }
    uint64_t batches() const { return batches_; }

  private:
    struct Arena {
        enum State { Filling, Sealed, Scoring } state = Filling;
        std::unique_ptr<Pinned<unsigned char>> E; // [cap][N][D] of the transport's element type
        std::unique_ptr<Pinned<float>> V, w, conf, cw, agree;
        std::unique_ptr<Pinned<int32_t>> winner, rank;
        std::vector<std::promise<score::completions::Tally>> promises;
        std::vector<std::optional<uint16_t>> failed;
        int reserved = 0;                      // slots handed out (under mu_)
        std::atomic<int> ready{0};             // slots completely written
        std::atomic<int> reserved_at_seal{-1}; // set by the worker when it takes the arena
        std::chrono::steady_clock::time_point first;
    };
    struct Bucket {
        int N = 0, D = 0, M = 0, cap = 0;
        Arena arena[2];
        int open = 0;
    };
    using Key = std::tuple<int, int, int>;

    Bucket &bucket_for(int N, int D, int M) { // under mu_
        auto it = buckets_.find(Key{N, D, M});
        if (it != buckets_.end()) return *it->second;
        auto b = std::make_unique<Bucket>();
        b->N = N, b->D = D, b->M = M;
        const size_t esize = opt_.transport == LWC_ELEM_F64 ? 8 : (opt_.transport == LWC_ELEM_BF16 ? 2 : 4);
        const size_t per = (size_t)N * std::max(D, 1) * esize;
        b->cap = (int)std::max<size_t>(1, std::min<size_t>((size_t)opt_.max_batch, opt_.max_arena_bytes / per));
        for (Arena &a : b->arena) { // pinned staging lives as long as the batcher: cudaHostAlloc costs milliseconds
            const size_t R = (size_t)b->cap, M1 = (size_t)std::max(M, 1);
            a.E = std::make_unique<Pinned<unsigned char>>(ctx_, R * per);
            a.V = std::make_unique<Pinned<float>>(ctx_, R * M1 * N);
            a.w = std::make_unique<Pinned<float>>(ctx_, R * M1);
            a.conf = std::make_unique<Pinned<float>>(ctx_, R * N);
            a.cw = std::make_unique<Pinned<float>>(ctx_, R * N);
            a.agree = std::make_unique<Pinned<float>>(ctx_, R * M1);
            a.winner = std::make_unique<Pinned<int32_t>>(ctx_, R);
            a.rank = std::make_unique<Pinned<int32_t>>(ctx_, R * N);
        }
        return *(buckets_[Key{N, D, M}] = std::move(b));
    }

    void run() {
        for (;;) {
            Bucket *b = nullptr;
            Arena *a = nullptr;
            {
                std::unique_lock<std::mutex> lk(mu_);
                for (;;) {
                    // a sealed arena first; else the partly filled arena that has lingered longest past its deadline
                    const auto now = std::chrono::steady_clock::now();
                    auto soonest = now + std::chrono::hours(1);
                    bool any = false;
                    for (auto &kv : buckets_)
                        for (Arena &x : kv.second->arena) {
                            if (x.reserved == 0 || x.state == Arena::Scoring) continue;
                            any = true;
                            const auto due = x.first + std::chrono::microseconds(opt_.max_wait_us);
                            if (x.state == Arena::Sealed || due <= now || stop_) {
                                if (!a || x.first < a->first) b = kv.second.get(), a = &x;
                            } else if (due < soonest) {
                                soonest = due;
                            }
                        }
                    if (a) break;
                    if (stop_ && !any) return;
                    if (any)
                        cv_work_.wait_until(lk, soonest);
                    else
                        cv_work_.wait(lk);
                }
                a->state = Arena::Scoring; // no further slots are handed out
                a->reserved_at_seal.store(a->reserved, std::memory_order_release);
                if (&b->arena[b->open] == a && b->arena[b->open ^ 1].state == Arena::Filling) b->open ^= 1;
                // submitters still writing their slots finish first
                cv_work_.wait(lk, [&] { return a->ready.load(std::memory_order_acquire) == a->reserved; });
            }
            score(*b, *a);
            {
                std::lock_guard<std::mutex> lk(mu_);
                a->promises.clear();
                a->failed.clear();
                a->reserved = 0;
                a->ready.store(0);
                a->reserved_at_seal.store(-1);
                a->state = Arena::Filling;
            }
            cv_space_.notify_all();
        }
    }

    void score(Bucket &b, Arena &a) {
        const int R = a.reserved, N = b.N, D = b.D, M = b.M;
        try {
            lwc_batch_desc d{};
            d.R = R, d.N = N, d.D = D, d.M = M;
            d.similarity = D ? LWC_SIM_COSINE : LWC_SIM_IDENTITY;
            d.precision = opt_.precision;
            d.memory = LWC_MEM_HOST;
            d.w_per_request = 1;
            ctx_.check(lwc_score_batch_typed(ctx_.get(), &d, opt_.transport, D ? a.E->data() : nullptr, a.V->data(), a.w->data(),
                                             a.conf->data(), a.cw->data(), a.agree->data(), a.winner->data(), a.rank->data(), nullptr));
            ++batches_;
            for (int r = 0; r < R; ++r) {
                score::completions::Tally t;
                t.weight.assign(a.cw->data() + (size_t)r * N, a.cw->data() + (size_t)(r + 1) * N);
                t.confidence.assign(a.conf->data() + (size_t)r * N, a.conf->data() + (size_t)(r + 1) * N);
                t.rank.assign(a.rank->data() + (size_t)r * N, a.rank->data() + (size_t)(r + 1) * N);
                t.winner = a.winner->data()[r];
                t.voter_confidence.resize(M);
                for (int m = 0; m < M; ++m) {
                    const float ag = a.agree->data()[(size_t)r * M + m];
                    if (!std::isnan(ag)) t.voter_confidence[m] = ag;
                }
                t.all_votes_failed = a.failed[r];
                a.promises[r].set_value(std::move(t));
            }
        } catch (...) {
            for (auto &p : a.promises) {
                try {
                    p.set_exception(std::current_exception());
                } catch (...) {
                }
            }
        }
    }

    Context ctx_;
    Options opt_;
    std::mutex mu_;
    std::condition_variable cv_work_, cv_space_;
    std::map<Key, std::unique_ptr<Bucket>> buckets_;
    bool stop_ = false;
    std::atomic<uint64_t> batches_{0};
    std::thread worker_;
};

} // namespace lwc
