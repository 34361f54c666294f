// test_host.cpp -- exercises llm-weighted-consensus_b200/host/lwc_host.hpp.
//   ./test_host            CPU box: type mirror + error mapping (no device -> ResponseError 503), K2 merge
//   ./test_host --gpu      B200: 8 threads submit 400 mixed-shape requests through the Batcher; every
//                          result must equal a single-request lwc_score_batch call bit for bit
#include <atomic>
#include <numeric>
#include <cstdio>
#include <random>

#include "../../llm-weighted-consensus_b200/host/lwc_host.hpp"

using namespace lwc;
using score::completions::Request;
using score::completions::Tally;
using score::completions::VoterChoice;

#define REQUIRE(c)                                                        \
    do {                                                                  \
        if (!(c)) {                                                       \
            std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);    \
            return 1;                                                     \
        }                                                                 \
    } while (0)

static int cpu_checks() {
    // client.rs:390-408 through the host types (golden K2)
    std::vector<VoterChoice> v(2);
    v[0].error_code = 429, v[1].error_code = 429;
    REQUIRE(merge_error_codes(v).value() == 429);
    v[0].error_code = 400;
    REQUIRE(merge_error_codes(v).value() == 400);
    v[1].error_code = 500;
    REQUIRE(merge_error_codes(v).value() == 500);
    v[1].error_code.reset(); // a voter without an error: not an all-failed request
    REQUIRE(!merge_error_codes(v).has_value());
    REQUIRE(merge_error_codes({}).value() == 500); // AllVotesFailed(None) -> 500 (error.rs:41)

    // streamed-chunk aggregation (response.rs:25-71,121-173,203-208) and the gather / write-back around the
    // tally (client.rs:384-456)
    namespace st = score::completions::response::streaming;
    st::ChatCompletionChunk agg;
    agg.id = "rsp";
    for (uint64_t i = 0; i < 2; ++i) { // the request's two candidate choices come first
        st::Choice c;
        c.index = i;
        c.delta.content = i == 0 ? "yes" : "no";
        agg.choices.push_back(c);
    }
    st::ChatCompletionChunk a, b, c3;
    {
        st::Choice c;
        c.index = 2; // voter 0, first piece
        c.delta.content = "{\"resp";
        c.model = "llm-a";
        c.model_index = 0;
        a.choices.push_back(c);
        c = st::Choice{};
        c.index = 2; // voter 0, last piece: vote and weight arrive at the end
        c.delta.content = "onse\": 1}";
        c.delta.vote = std::vector<double>{0.25, 0.75};
        c.weight = 2.0;
        c.model = "ignored: first value wins";
        c.finish_reason = "stop";
        b.choices.push_back(c);
        c = st::Choice{};
        c.index = 3; // voter 1 failed
        c.error = error::ResponseError(429, "rate limited");
        c.weight = 1.0;
        c3.choices.push_back(c);
    }
    agg.push(a);
    agg.push(b);
    agg.push(c3);
    REQUIRE(agg.choices.size() == 4);
    REQUIRE(*agg.choices[2].delta.content == "{\"response\": 1}");
    REQUIRE(*agg.choices[2].model == "llm-a" && *agg.choices[2].model_index == 0);
    REQUIRE(*agg.choices[2].finish_reason == "stop");
    st::ChatCompletionChunk late;
    {
        st::Choice c;
        c.index = 2;
        c.delta.vote = std::vector<double>{1.0, 0.0}; // a second vote must not replace the first
        c.weight = 9.0;
        late.choices.push_back(c);
    }
    agg.push(late);
    REQUIRE((*agg.choices[2].delta.vote)[1] == 0.75 && *agg.choices[2].weight == 2.0);
    const Request rq = score::completions::request_from_aggregate(agg, 2);
    REQUIRE(rq.n_choices == 2 && rq.voters.size() == 2);
    REQUIRE(rq.voters[0].vote.has_value() && (*rq.voters[0].vote)[0] == 0.25 && *rq.voters[0].weight == 2.0);
    REQUIRE(!rq.voters[0].error_code.has_value());
    REQUIRE(!rq.voters[1].vote.has_value() && *rq.voters[1].error_code == 429);
    REQUIRE(!merge_error_codes(rq.voters).has_value()); // one voter succeeded
    Tally t; // what the reference arithmetic gives for this aggregate: t = 2 * (0.25, 0.75)
    t.weight = {0.5, 1.5};
    t.confidence = {0.25, 0.75};
    t.voter_confidence = {0.25 * 0.25 + 0.75 * 0.75, std::nullopt};
    score::completions::apply_tally(agg, t, 2);
    REQUIRE(*agg.choices[0].weight == 0.5 && *agg.choices[1].confidence == 0.75);
    REQUIRE(std::fabs(*agg.choices[2].confidence - 0.625) < 1e-12);
    REQUIRE(!agg.choices[3].confidence.has_value()); // no vote: untouched (client.rs:440)
    for (const auto &ch : agg.choices)
        REQUIRE(!ch.delta.content && !ch.delta.vote && !ch.finish_reason && !ch.error); // client.rs:451-454
    // ---- candidate sources: a multichat completion's generations become the candidates of a score request, and the
    // candidate order (= row order of E) is the request's choice order (request.rs:84-89, client.rs:144-170,1145-1152)
    {
        namespace rq = score::completions::request;
        multichat::completions::response::unary::ChatCompletion mc;
        mc.id = "mc-1";
        for (uint64_t i = 0; i < 4; ++i) {
            multichat::completions::response::unary::Choice c;
            c.index = 3 - i; // stored out of order: lookup is by `index`, not by position
            c.message.content = "answer " + std::to_string(3 - i);
            if (3 - i == 1) c.message.reasoning = "because";
            if (3 - i == 2) c.error = error::ResponseError(500, "generation failed");
            mc.choices.push_back(c);
        }
        rq::MultichatArchive archive{{mc.id, mc}};
        std::vector<rq::Choice> choices = rq::choices_from_multichat(mc); // failed generation 2 is left out
        REQUIRE(choices.size() == 3);
        choices.push_back(rq::Choice::Text("a literal candidate"));
        const auto internal = rq::convert_choices_to_internal_choices(archive, choices);
        const auto texts = rq::choice_texts(internal);
        REQUIRE(texts.size() == 4);
        REQUIRE(texts[0] == "answer 3" && texts[1] == "because\n\nanswer 1" && texts[2] == "answer 0");
        REQUIRE(texts[3] == "a literal candidate");
        REQUIRE(internal[1].multichat_choice.index == 1);
        try {
            rq::convert_choices_to_internal_choices(archive, {rq::Choice::MultichatCompletion("mc-1", 9)});
            REQUIRE(false);
        } catch (const error::ResponseError &e) { // Error::InvalidCompletionChoiceIndex -> 400 (error.rs:43)
            REQUIRE(e.code == 400 && e.message.find("invalid choice_index for completion mc-1: 9") != std::string::npos);
        }
    }
    // ---- the host half of get_vote's numeric tail (client.rs:1766-1784): which alternatives reach the device
    {
        using chat::completions::response::TopLogprob;
        std::vector<TopLogprob> alts = {{"`B", -0.1}, {"`A", -2.5}, {"`", -3.0}, {"`Z", -1.0}, {"`C", std::nullopt}, {"\xc3\xa9" "A", -0.5}};
        auto leaf_of = [](char c) -> std::optional<int32_t> { // lowest branch holds the leaves A, B, C -> choices 4, 7, 2
            if (c == 'A') return 4;
            if (c == 'B') return 7;
            if (c == 'C') return 2;
            return std::nullopt;
        };
        float lp[8];
        int32_t leaf[8];
        score::completions::fill_vote_pairs(alts, 1, leaf_of, 8, lp, leaf);
        REQUIRE(leaf[0] == 7 && lp[0] == -0.1f);
        REQUIRE(leaf[1] == 4 && lp[1] == -2.5f);
        REQUIRE(leaf[2] == -1);            // token too short
        REQUIRE(leaf[3] == -1);            // 'Z' is no sibling leaf
        REQUIRE(leaf[4] == -1);            // logprob None
        REQUIRE(leaf[5] == -1);            // byte 1 is inside a multi-byte char
        REQUIRE(leaf[6] == -1 && leaf[7] == -1 && std::isnan(lp[7])); // padding
    }
    // ---- weight::Fetchers dispatch (weight.rs:45-63) with the static fetcher (weight.rs:76-97)
    {
        namespace wt = score::completions::weight;
        struct Ctx {};
        struct Req {
            std::string prompt;
        };
        score::model::Model model;
        const double ws[3] = {1.0, 0.25, 2.5};
        for (size_t i = 0; i < 3; ++i) {
            score::model::Llm l;
            l.id = "llm" + std::to_string(i);
            l.index = i;
            l.weight.weight = ws[i];
            REQUIRE(!l.weight.validate().has_value());
            model.llms.push_back(l);
        }
        wt::Fetchers<Ctx, Req> fetchers;
        fetchers.static_ = std::make_shared<wt::StaticFetcher<Ctx, Req>>();
        auto got = fetchers.fetch(Ctx{}, std::make_shared<const Req>(Req{"p"}), model);
        REQUIRE(got.second.type == score::WeightType::Static);
        REQUIRE(got.first.size() == 3 && got.first[0] == 1.0 && got.first[1] == 0.25 && got.first[2] == 2.5);
        score::llm::Weight bad;
        bad.weight = 0.0; // llm/mod.rs:650-660
        REQUIRE(bad.validate().has_value());
        bad = score::llm::Weight{};
        bad.type = score::WeightType::TrainingTable;
        bad.base_weight = 1.0, bad.min_weight = 2.0, bad.max_weight = 3.0; // base < min, llm/mod.rs:671-687
        REQUIRE(bad.validate().has_value());
    }
    return 0;
}

static int no_device_check() {
    try {
        Context c(0);
    } catch (const error::ResponseError &e) {
        REQUIRE(e.code == 503);
        REQUIRE(e.message.find("no CPU path") != std::string::npos);
        std::printf("no device: %s\n", e.what());
        return 0;
    }
    std::printf("a CUDA device is present; skipping the no-device check\n");
    return 0;
}

static int gpu_checks() {
    Batcher::Options opt;
    opt.max_wait_us = 500;
    Batcher batcher(0, opt);
    const int shapes[3][3] = {{16, 1536, 4}, {5, 64, 3}, {32, 256, 7}};
    const int T = 8, PER = 50;
    std::vector<std::vector<Request>> reqs(T);
    std::vector<std::vector<embeddings::response::CreateEmbeddingResponse>> embs(T);
    std::vector<std::vector<Tally>> got(T);
    for (int t = 0; t < T; ++t) {
        std::mt19937 g(100 + t);
        std::normal_distribution<double> nd;
        std::uniform_real_distribution<double> ud(0.0, 1.0);
        embs[t].resize(PER);
        reqs[t].resize(PER);
        for (int i = 0; i < PER; ++i) {
            const int *s = shapes[(t + i) % 3];
            const int N = s[0], D = s[1], M = s[2];
            auto &er = embs[t][i];
            er.data.resize(N);
            for (int n = 0; n < N; ++n) {
                er.data[n].index = (uint64_t)((n * 7 + 3) % N); // scrambled arrival order, as over the wire
                if (std::gcd(7, N) != 1) er.data[n].index = n;
                er.data[n].embedding.resize(D);
                for (auto &x : er.data[n].embedding) x = nd(g);
            }
            Request &rq = reqs[t][i];
            rq.n_choices = N;
            rq.embeddings = (i % 5 == 4) ? nullptr : &er; // some pure-tally requests (similarity = identity)
            rq.voters.resize(M);
            for (int m = 0; m < M; ++m) {
                if (ud(g) < 0.1) continue; // voter without a vote
                std::vector<double> vote(N);
                double sum = 0;
                for (auto &x : vote) sum += (x = ud(g));
                for (auto &x : vote) x /= sum;
                rq.voters[m].vote = vote;
                if (ud(g) > 0.05) rq.voters[m].weight = 0.5 + 1.5 * ud(g);
            }
        }
    }
    std::vector<std::thread> th;
    std::atomic<int> failures{0};
    for (int t = 0; t < T; ++t)
        th.emplace_back([&, t] {
            std::vector<std::future<Tally>> f;
            for (auto &rq : reqs[t]) f.push_back(batcher.submit(rq));
            for (auto &x : f) {
                try {
                    got[t].push_back(x.get());
                } catch (const std::exception &e) {
                    std::printf("request failed: %s\n", e.what());
                    ++failures;
                }
            }
        });
    for (auto &x : th) x.join();
    REQUIRE(failures == 0);
    std::printf("batches launched: %llu for %d requests\n", (unsigned long long)batcher.batches(), T * PER);
    REQUIRE(batcher.batches() < (uint64_t)T * PER); // requests were actually batched

    // every batched result == the same request alone through the Batcher's own context
    Batcher solo(0);
    for (int t = 0; t < T; ++t)
        for (int i = 0; i < PER; ++i) {
            Tally a = solo.submit(reqs[t][i]).get();
            const Tally &b = got[t][i];
            REQUIRE(a.winner == b.winner);
            REQUIRE(a.rank == b.rank);
            REQUIRE(a.weight.size() == b.weight.size());
            for (size_t k = 0; k < a.weight.size(); ++k) {
                REQUIRE(a.weight[k] == b.weight[k]);
                REQUIRE(a.confidence[k] == b.confidence[k]);
            }
            for (size_t m = 0; m < a.voter_confidence.size(); ++m) {
                REQUIRE(a.voter_confidence[m].has_value() == b.voter_confidence[m].has_value());
                REQUIRE(a.voter_confidence[m].has_value() == reqs[t][i].voters[m].vote.has_value());
                if (a.voter_confidence[m]) REQUIRE(*a.voter_confidence[m] == *b.voter_confidence[m]);
            }
            double s = 0;
            for (double c : b.confidence) s += c;
            if (b.winner >= 0) {
                REQUIRE(std::fabs(s - 1.0) < 1e-5);
                REQUIRE(b.rank[0] == b.winner);
            }
        }
    // the element type on the wire is a batcher option: f64 (rounded on the device) must give the fp32 transport's
    // bits, bf16 (opt-in) must stay within its stated 1e-2 of the top confidence
    {
        Batcher::Options o64 = opt, o16 = opt;
        o64.transport = LWC_ELEM_F64;
        o16.transport = LWC_ELEM_BF16;
        Batcher b64(0, o64), b16(0, o16);
        for (int t = 0; t < 2; ++t)
            for (int i = 0; i < 30; ++i) {
                if (!reqs[t][i].embeddings) continue;
                const Tally &a = got[t][i];
                const Tally x = b64.submit(reqs[t][i]).get(), y = b16.submit(reqs[t][i]).get();
                REQUIRE(x.winner == a.winner && x.rank == a.rank);
                double top = 0;
                for (double c : a.confidence) top = std::max(top, std::fabs(c));
                for (size_t k = 0; k < a.confidence.size(); ++k) {
                    REQUIRE(x.confidence[k] == a.confidence[k] && x.weight[k] == a.weight[k]);
                    REQUIRE(std::fabs(y.confidence[k] - a.confidence[k]) <= 1e-2 * top);
                }
            }
    }
    // from streamed chunks to the written-back aggregate (client.rs:362, 384-456) with the tally on the GPU:
    // similarity = identity, so the expected numbers are the reference's own (golden K-style case)
    {
        namespace st = score::completions::response::streaming;
        st::ChatCompletionChunk agg;
        for (uint64_t i = 0; i < 3; ++i) {
            st::Choice c;
            c.index = i;
            agg.choices.push_back(c);
        }
        const double votes[3][3] = {{0.5, 0.25, 0.25}, {0.0, 1.0, 0.0}, {0.125, 0.125, 0.75}};
        const double weights[3] = {1.0, 2.0, 4.0};
        for (int m = 0; m < 3; ++m) {
            st::ChatCompletionChunk piece, last;
            st::Choice c;
            c.index = 3 + m;
            c.delta.content = "thinking";
            piece.choices.push_back(c);
            c = st::Choice{};
            c.index = 3 + m;
            c.delta.vote = std::vector<double>(votes[m], votes[m] + 3);
            c.weight = weights[m];
            last.choices.push_back(c);
            agg.push(piece);
            agg.push(last);
        }
        const Request rq = score::completions::request_from_aggregate(agg, 3);
        const Tally tl = batcher.submit(rq).get();
        score::completions::apply_tally(agg, tl, 3);
        const double cw[3] = {0.5 * 1 + 0.125 * 4, 0.25 * 1 + 1.0 * 2 + 0.125 * 4, 0.25 * 1 + 0.75 * 4};
        const double sum = cw[0] + cw[1] + cw[2];
        for (int j = 0; j < 3; ++j) {
            REQUIRE(std::fabs(*agg.choices[j].weight - cw[j]) < 1e-6);
            REQUIRE(std::fabs(*agg.choices[j].confidence - cw[j] / sum) < 1e-6);
        }
        for (int m = 0; m < 3; ++m) {
            double a = 0;
            for (int j = 0; j < 3; ++j) a += cw[j] / sum * votes[m][j];
            REQUIRE(std::fabs(*agg.choices[3 + m].confidence - a) < 1e-6);
            REQUIRE(!agg.choices[3 + m].delta.vote.has_value());
        }
        REQUIRE(tl.winner == 2 && tl.rank[0] == 2);
    }
    // training-table weights through the host wrapper: the prompt equal to table row 5 has row 5 as its first
    // neighbour with cosine 1, and an all-equal quality table returns the base weights
    {
        const int rows = 64, D = 32, M = 3;
        std::mt19937 g(9);
        std::normal_distribution<float> nd;
        std::vector<float> T((size_t)rows * D), Q((size_t)rows * M, 0.5f), q(D);
        for (auto &x : T) x = nd(g);
        for (int k = 0; k < D; ++k) q[k] = 2.0f * T[5 * D + k];
        const Context &c = batcher.context();
        c.table_load(T, Q, rows, D, M);
        Context::TableLimits lim{{1.0f, 2.0f, 3.0f}, {0.5f, 1.0f, 1.5f}, {2.0f, 4.0f, 6.0f}};
        std::vector<int32_t> idx;
        std::vector<float> sim;
        const std::vector<float> w = c.table_weights(q, 1, 4, lim, &idx, &sim);
        REQUIRE(idx[0] == 5 && std::fabs(sim[0] - 1.0f) < 1e-6f);
        for (int m = 0; m < M; ++m) REQUIRE(std::fabs(w[m] - lim.base[m]) < 1e-6f);
        try {
            lim.min[0] = 5.0f; // min > base: the reference's validate() refuses it (llm/mod.rs:671-687)
            c.table_weights(q, 1, 4, lim);
            REQUIRE(false);
        } catch (const error::ResponseError &e) {
            REQUIRE(e.code == 400);
        }
    }
    // weight::Fetcher for WeightType::TrainingTable: 24 threads fetch at once; the worker batches their prompts into a
    // few lwc_table_weights calls; every result equals the direct host call for the same prompt
    {
        namespace wt = score::completions::weight;
        struct Ctx {};
        struct Req {
            std::vector<double> prompt_embedding; // what the embeddings service would return for the prompt
        };
        const int rows = 512, D = 64, M = 4, T2 = 24;
        std::mt19937 g(17);
        std::normal_distribution<float> nd;
        std::uniform_real_distribution<float> ud(0.f, 1.f);
        std::vector<float> Tt((size_t)rows * D), Q((size_t)rows * M);
        for (auto &x : Tt) x = nd(g);
        for (auto &x : Q) x = ud(g);
        const Context &c = batcher.context();
        c.table_load(Tt, Q, rows, D, M);
        score::model::Model model;
        model.weight.type = score::WeightType::TrainingTable;
        model.weight.top = 8;
        Context::TableLimits lim;
        for (int m = 0; m < M; ++m) {
            score::model::Llm l;
            l.index = (size_t)m;
            l.weight.type = score::WeightType::TrainingTable;
            l.weight.base_weight = 1.0 + 0.5 * m, l.weight.min_weight = 0.5, l.weight.max_weight = 4.0;
            model.llms.push_back(l);
            lim.base.push_back((float)l.weight.base_weight), lim.min.push_back(0.5f), lim.max.push_back(4.0f);
        }
        std::atomic<int> embed_calls{0};
        auto tt = std::make_shared<wt::B200TrainingTableFetcher<Ctx, Req>>(
            c, D, [&](Ctx, const Req &r, const score::model::Model &) {
                ++embed_calls;
                embeddings::response::CreateEmbeddingResponse er;
                er.data.resize(1);
                er.data[0].embedding = r.prompt_embedding;
                return er;
            });
        wt::Fetchers<Ctx, Req> fetchers;
        fetchers.static_ = std::make_shared<wt::StaticFetcher<Ctx, Req>>();
        fetchers.training_table = tt;
        std::vector<std::shared_ptr<const Req>> rq2;
        for (int t = 0; t < T2; ++t) {
            Req r;
            r.prompt_embedding.resize(D);
            for (auto &x : r.prompt_embedding) x = nd(g);
            rq2.push_back(std::make_shared<const Req>(std::move(r)));
        }
        std::vector<std::pair<std::vector<double>, wt::Data>> res(T2);
        std::vector<std::thread> th2;
        for (int t = 0; t < T2; ++t) th2.emplace_back([&, t] { res[t] = fetchers.fetch(Ctx{}, rq2[t], model); });
        for (auto &x : th2) x.join();
        REQUIRE(embed_calls == T2);
        std::printf("table-weight calls: %llu for %d fetches\n", (unsigned long long)tt->calls(), T2);
        REQUIRE(tt->calls() >= 1 && tt->calls() <= (uint64_t)T2);
        for (int t = 0; t < T2; ++t) {
            REQUIRE(res[t].second.type == score::WeightType::TrainingTable);
            REQUIRE(res[t].second.training_table.embeddings_response.data[0].embedding == rq2[t]->prompt_embedding);
            std::vector<float> q(rq2[t]->prompt_embedding.begin(), rq2[t]->prompt_embedding.end());
            const std::vector<float> w = c.table_weights(q, 1, 8, lim);
            REQUIRE(res[t].first.size() == (size_t)M);
            for (int m = 0; m < M; ++m) REQUIRE((float)res[t].first[m] == w[m]);
        }
        // a model whose limits the reference's validate() refuses: the error arrives through the future as 400
        score::model::Model badm = model;
        badm.llms[0].weight.min_weight = 9.0;
        try {
            tt->fetch(Ctx{}, rq2[0], badm).get();
            REQUIRE(false);
        } catch (const error::ResponseError &e) {
            REQUIRE(e.code == 400);
        }
    }
    // argument errors surface as ResponseError 400, like Error::ExpectedTwoOrMoreChoices would
    try {
        Request bad;
        bad.n_choices = 0;
        batcher.submit(bad);
        REQUIRE(false);
    } catch (const error::ResponseError &e) {
        REQUIRE(e.code == 400);
    }
    return 0;
}

int main(int argc, char **argv) {
    const bool gpu = argc > 1 && std::string(argv[1]) == "--gpu";
    if (cpu_checks()) return 1;
    if (gpu ? gpu_checks() : no_device_check()) return 1;
    std::printf("OK\n");
    return 0;
}
