// e2e_host.cpp -- end to end FROM THE REFERENCE'S OWN TYPES, wall clock.
//
// What a Rust caller of the reference holds when the scoring path starts is not a contiguous fp32 batch: it is,
// per request, a CreateEmbeddingResponse whose rows are separately allocated Vec<f64> in arrival order
// (src/embeddings/response.rs:4-24) and, per voter, an Option<Vec<Decimal>> vote and an Option<Decimal> weight
// (src/score/completions/response.rs:86-107,184-190).  This tool builds BASELINE.json's config 2 in exactly that
// shape (lwc::embeddings::response::*, lwc::score::completions::Request), lets T host threads submit the
// requests to lwc::Batcher -- which converts f64 -> fp32 straight into pinned [R][N][D] arenas, calls
// lwc_score_batch and hands back Tally futures -- and times whole rounds with the host clock.  Everything the
// real boundary pays is inside the timed region: conversion, row placement by Embedding.index, batching, H2D,
// kernels, D2H, Tally construction.  bench.py runs it and puts the JSON line under `e2e_from_reference_types`.
//
//   tools/e2e_host [--device 0] [--threads T] [--requests 4096] [--rounds 5] [--n 16] [--d 1536] [--m 4] [--batch 512] [--transport fp32|f64|bf16]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <random>
#include <string>

#include "../llm-weighted-consensus_b200/host/lwc_host.hpp"

using namespace lwc;
using score::completions::Request;
using score::completions::Tally;

int main(int argc, char **argv) {
    int device = 0, R = 4096, rounds = 5, N = 16, D = 1536, M = 4, batch = 512, transport = LWC_ELEM_F32;
    int T = (int)std::min(32u, std::max(1u, std::thread::hardware_concurrency()));
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string k = argv[i];
        const int v = std::atoi(argv[i + 1]);
        if (k == "--transport") {
            const std::string t = argv[i + 1];
            transport = t == "f64" ? LWC_ELEM_F64 : (t == "bf16" ? LWC_ELEM_BF16 : LWC_ELEM_F32);
            continue;
        }
        if (k == "--device") device = v;
        else if (k == "--threads") T = v;
        else if (k == "--requests") R = v;
        else if (k == "--rounds") rounds = v;
        else if (k == "--n") N = v;
        else if (k == "--d") D = v;
        else if (k == "--m") M = v;
        else if (k == "--batch") batch = v;
    }
    try {
        // ---- the batch as the reference would hold it: heap rows of f64, scrambled arrival order
        std::vector<embeddings::response::CreateEmbeddingResponse> embs((size_t)R);
        std::vector<Request> reqs((size_t)R);
        {
            std::vector<std::thread> gen;
            for (int t = 0; t < T; ++t)
                gen.emplace_back([&, t] {
                    std::mt19937_64 g(1000 + t);
                    std::normal_distribution<double> nd;
                    std::uniform_real_distribution<double> ud(0.0, 1.0);
                    for (int r = t; r < R; r += T) {
                        auto &er = embs[(size_t)r];
                        er.data.resize((size_t)N);
                        std::vector<double> centre((size_t)D);
                        for (auto &x : centre) x = nd(g);
                        for (int n = 0; n < N; ++n) {
                            er.data[(size_t)n].index = (uint64_t)(std::gcd(7, N) == 1 ? (n * 7 + 3) % N : n); // arrival order != row order
                            auto &row = er.data[(size_t)n].embedding;
                            row.resize((size_t)D);
                            for (int k = 0; k < D; ++k) row[(size_t)k] = centre[(size_t)k] * (n % 4 == 0 ? 1.0 : 0.3) + 0.35 * nd(g);
                        }
                        Request &rq = reqs[(size_t)r];
                        rq.n_choices = N;
                        rq.embeddings = &er;
                        rq.voters.resize((size_t)M);
                        for (int m = 0; m < M; ++m) {
                            if (ud(g) < 0.02) continue; // failed voter: no vote (client.rs:410)
                            std::vector<double> vote((size_t)N, 0.0);
                            if (ud(g) < 0.5) {
                                vote[(size_t)(g() % N)] = 1.0; // one-hot (client.rs:1796-1798)
                            } else {
                                double s = 0;
                                for (auto &x : vote) s += (x = ud(g));
                                for (auto &x : vote) x /= s;
                            }
                            rq.voters[(size_t)m].vote = std::move(vote);
                            rq.voters[(size_t)m].weight = 0.5 + 0.01 * (double)(g() % 151);
                        }
                    }
                });
            for (auto &x : gen) x.join();
        }
        Batcher::Options opt;
        opt.max_batch = batch; // two arenas per shape: the submitters fill one while the GPU works through the other
        opt.max_wait_us = 300;
        opt.transport = transport;
        opt.max_arena_bytes = (size_t)1 << 30;
        Batcher batcher(device, opt);
        std::vector<Tally> out((size_t)R);
        auto round = [&] {
            std::vector<std::thread> th;
            for (int t = 0; t < T; ++t)
                th.emplace_back([&, t] {
                    // contiguous shares: thread t owns requests [b, e)
                    const int b = (int)((long long)R * t / T), e = (int)((long long)R * (t + 1) / T);
                    std::vector<std::future<Tally>> f;
                    f.reserve((size_t)(e - b));
                    for (int r = b; r < e; ++r) f.push_back(batcher.submit(reqs[(size_t)r]));
                    for (int r = b; r < e; ++r) out[(size_t)r] = f[(size_t)(r - b)].get();
                });
            for (auto &x : th) x.join();
        };
        round(); // warm-up: pinned arenas, device buffers, kernel attributes
        round();
        const uint64_t batches0 = batcher.batches();
        std::vector<double> ms;
        const auto t_all = std::chrono::steady_clock::now();
        for (int k = 0; k < rounds; ++k) {
            const auto t0 = std::chrono::steady_clock::now();
            round();
            ms.push_back(std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
        }
        const double total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_all).count();
        // What bounds this leg: the same T threads doing ONLY the pass every host implementation owes these inputs --
        // read each f64 heap row once, write it once as fp32 (lwc_stage_f64 into a plain arena; no batching, no GPU).
        double convert_only = 0;
        if (transport == LWC_ELEM_F32) {
            std::vector<float> arena((size_t)R * N * D);
            lwc_ctx *cx = batcher.context().get();
            auto pass = [&] {
                std::vector<std::thread> th;
                for (int t = 0; t < T; ++t)
                    th.emplace_back([&, t] {
                        const int b = (int)((long long)R * t / T), e = (int)((long long)R * (t + 1) / T);
                        for (int r = b; r < e; ++r)
                            for (const auto &row : embs[(size_t)r].data)
                                lwc_stage_f64(cx, row.embedding.data(), (size_t)D, arena.data() + ((size_t)r * N + row.index) * D);
                    });
                for (auto &x : th) x.join();
            };
            pass();
            const auto c0 = std::chrono::steady_clock::now();
            for (int k = 0; k < rounds; ++k) pass();
            convert_only = (double)R * rounds / std::chrono::duration<double>(std::chrono::steady_clock::now() - c0).count();
        }
        // sanity of what came back (the parity tests are elsewhere): confidences of scored requests sum to 1
        int scored = 0, bad = 0;
        for (const Tally &t : out) {
            if (t.winner < 0) continue;
            ++scored;
            double s = 0;
            for (double c : t.confidence) s += c;
            if (std::fabs(s - 1.0) > 1e-4 || t.rank.empty() || t.rank[0] != t.winner) ++bad;
        }
        std::sort(ms.begin(), ms.end());
        std::printf("{\"value\": %.1f, \"unit\": \"requests/s\", \"ms_per_round\": %.3f, \"ms_per_round_min\": %.3f, \"rounds\": %d, "
                    "\"requests_per_round\": %d, \"submit_threads\": %d, \"max_batch\": %d, \"transport\": \"%s\", \"batches_per_round\": %.2f, \"scored\": %d, \"inconsistent\": %d, "
                    "\"host_bytes_converted_per_round\": %.0f, \"convert_only_requests_per_s\": %.1f, \"timing\": \"host wall clock around whole rounds\", "
                    "\"api\": \"lwc::Batcher::submit(Request{CreateEmbeddingResponse of Vec<f64> rows, Option votes}) -> future<Tally>; "
                    "f64->fp32 conversion, row placement, batching, H2D, kernels, D2H inside the timed region\", "
                    "\"shape\": \"%d requests x %d candidates x %d-dim f64, %d voters\"}\n",
                    (double)R * rounds / total_s, 1e3 * total_s / rounds, ms.front(), rounds, R, T, batch, transport == LWC_ELEM_F64 ? "f64 over PCIe, rounded on the device" : (transport == LWC_ELEM_BF16 ? "bf16 over PCIe (1e-2 tolerance)" : "fp32 over PCIe, converted by the submitting threads"),
                    (double)(batcher.batches() - batches0) / rounds, scored, bad, (double)R * N * D * 8.0, convert_only, R, N, D, M);
        return bad ? 2 : 0;
    } catch (const error::ResponseError &e) {
        std::fprintf(stderr, "e2e_host: %s\n", e.what());
        return 1;
    }
}
