// ff_stage.h -- the "queryExecution stage" of this path: lowers a netsDB Computation graph built from
// the FF UDFs to calls into the CUDA library.  It stands where PipelineStage::runPipeline would probe the
// join, run the projection lambda per pair and push partial blocks into the aggregation sink
// (queryExecution/source/PipelineStage.cc:334-930, lambdas/headers/JoinTuple.h:381-565,
// queryExecution/headers/CombinedShuffleSink.h:78-159).  Recognised shapes:
//
//   FFAggMatrix(FFTransposeMult|FFInputLayerJoin(a, b))                          -> one block GEMM
//   FFReluBiasSum(FFAggMatrix(join), bias) / FFTransposeBiasSum(..., bias)       -> the same GEMM with the bias join fused
//   FFOutputLayer(x, FFRowAggregate(x))                                          -> row softmax
//   LSTMThreeWaySum(x, y, z) / LSTMTwoSum(x, y)                                  -> one element-wise kernel
//   FFAggMatrix(LSTMHiddenState(o, c))                                           -> tanh map + block GEMM
//   FullyConnectedNetwork(x) / SemanticClassifier(x)                             -> nb200_udf_dense2 (two fp64 GEMMs)
//   the whole graph of ff::inference_unit (SimpleFF.cc:340-389)                  -> the fused stage with the
//                                                                                   label-exact precision cascade
// Anything else is rejected: there is no CPU fallback.
#pragma once
#include <mutex>
#include <set>
#include <sstream>
#include <thread>

#include "conv_stage.h"
#include "la_udfs.h"
#include "lstm_udfs.h"
#include "net_udfs.h"

namespace nb200_host {

inline int precision_from_env() {
    const char *e = getenv("NB200_PRECISION");
    if (!e) return NB200_PREC_BF16;
    if (!strcmp(e, "bf16x3")) return NB200_PREC_BF16X3;
    if (!strcmp(e, "fp64")) return NB200_PREC_FP64;
    if (!strcmp(e, "f16")) return NB200_PREC_F16;
    return NB200_PREC_BF16;
}

// ---- several GPUs from one client process (NB200_DEVICES=n) -------------------------------------------------------
// The reference partitions a set over the workers of the cluster (CombinedShuffleSink.h:95-98); here the batch rows of
// the sets named in NB200_SHARD_SETS (default "inputs") are split contiguously over n devices and every other set is
// replicated, so that each device runs the whole graph on its rows with no exchange.  The (db, set)
// namespace of the library is per device: the same lowering code runs once per device, on its own thread.
inline int &num_devices() { static int n = 1; return n; }
inline std::set<std::pair<std::string, std::string>> &sharded_sets() { static std::set<std::pair<std::string, std::string>> s; return s; }
inline bool shard_by_name(const std::string &set) {
    static std::set<std::string> names = [] {
        std::set<std::string> n;
        const char *e = getenv("NB200_SHARD_SETS");
        std::stringstream ss(e ? e : "inputs");
        std::string t;
        while (std::getline(ss, t, ',')) if (!t.empty()) n.insert(t);
        return n;
    }();
    return names.count(set) != 0;
}
// contiguous row shard of device d: sizes differ by at most one (netsdb_b200/shard.py row_shard)
inline void row_shard(int total, int world, int d, int *row0, int *rows) {
    const int base = total / world, extra = total % world;
    *rows = base + (d < extra ? 1 : 0);
    *row0 = d * base + std::min(d, extra);
}
inline void die_on(int st) {
    if (st != NB200_OK) {
        std::cerr << nb200_last_error() << std::endl;  // reference convention: print and exit(1)
        exit(1);
    }
}

inline void ensure_init() {
    static bool done = false;
    if (done) return;
    const char *d = getenv("NB200_DEVICE");
    const char *nd = getenv("NB200_DEVICES");
    const int n = nd ? atoi(nd) : 1;
    if (n > 1) {
        std::vector<int> devs;
        for (int i = 0; i < n; i++) devs.push_back(i);
        die_on(nb200_init_devices(devs.data(), n));
        num_devices() = n;
    } else {
        die_on(nb200_init(d ? atoi(d) : 0));
    }
    done = true;
}

// sendData for one FFMatrixBlock: the device set is created from the first block's meta
// total_rows / total_cols: extents to use when the block's meta leaves them at -1 (LSTMTest.cc:63-110 builds its blocks
// with the 4-argument constructor); the caller derives them from the block grid, i.e. the set is carried with its
// padding, exactly as the reference carries it
template <class Block>
inline int ship_block_one(const std::string &db, const std::string &set, pdb::Handle<Block> &b, int total_rows, int total_cols);
template <class Block>
inline int ship_block(const std::string &db, const std::string &set, pdb::Handle<Block> &b, int total_rows = -1, int total_cols = -1) {
    const int n = num_devices();
    if (n == 1) return ship_block_one(db, set, b, total_rows, total_cols);
    int st = NB200_OK;
    if (!shard_by_name(set)) {                       // replicated: the same block to every device
        for (int d = 0; d < n && st == NB200_OK; d++) {
            st = nb200_use_device(d);
            if (st == NB200_OK) st = ship_block_one(db, set, b, total_rows, total_cols);
        }
    } else {                                         // row-sharded: each device gets its rows of the block
        const int tr = b->getTotalRowNums() > 0 ? b->getTotalRowNums() : total_rows;
        const int tc = b->getTotalColNums() > 0 ? b->getTotalColNums() : total_cols;
        const int R = b->getRowNums(), C = b->getColNums();
        const int blo = b->getBlockRowIndex() * R, bhi = std::min(blo + R, tr);
        sharded_sets().insert(std::make_pair(db, set));
        for (int d = 0; d < n && st == NB200_OK; d++) {
            int r0, nr;
            row_shard(tr, n, d, &r0, &nr);
            const int lo = std::max(blo, r0), hi = std::min(bhi, r0 + nr);
            st = nb200_use_device(d);
            if (st != NB200_OK) break;
            nb200_set *s = nullptr;
            if (nb200_set_lookup(db.c_str(), set.c_str(), &s) != NB200_OK) st = nb200_set_create(db.c_str(), set.c_str(), nr, tc, R, C, &s);
            if (st == NB200_OK && hi > lo)
                st = nb200_upload_rect(s, lo - r0, b->getBlockColIndex() * C, b->getRawDataHandle()->c_ptr() + static_cast<size_t>(lo - blo) * C,
                                       hi - lo, C, C);
        }
    }
    nb200_use_device(0);
    return st;
}
template <class Block>
inline int ship_block_one(const std::string &db, const std::string &set, pdb::Handle<Block> &b, int total_rows, int total_cols) {
    nb200_set *s = nullptr;
    if (nb200_set_lookup(db.c_str(), set.c_str(), &s) != NB200_OK) {
        // nominal block size = size of the first block sent (block (0,0) in ff::loadMatrix's order)
        // FFMatrixMeta leaves unknown totals at -1, MatrixMeta at 0
        const int tr = b->getTotalRowNums() > 0 ? b->getTotalRowNums() : total_rows;
        const int tc = b->getTotalColNums() > 0 ? b->getTotalColNums() : total_cols;
        int st = nb200_set_create(db.c_str(), set.c_str(), tr, tc, b->getRowNums(),
                                  b->getColNums(), &s);
        if (st != NB200_OK) return st;
    }
    return nb200_upload_block(s, b->getBlockRowIndex(), b->getBlockColIndex(), b->getRawDataHandle()->c_ptr(), b->getRowNums(),
                              b->getColNums());
}

// getSetIterator: re-wrap one device block as an FFMatrixBlock (full-size, zero padded, reference meta)
template <class Block>
inline pdb::Handle<Block> make_block(int br, int bc, int rows, int cols, int tr, int tc) { return pdb::makeObject<Block>(br, bc, rows, cols, tr, tc); }
template <>
inline pdb::Handle<FFMatrixBlock> make_block<FFMatrixBlock>(int br, int bc, int rows, int cols, int tr, int tc) {
    return pdb::makeObject<FFMatrixBlock>(br, bc, rows, cols, tr, tc, false);
}
template <class Block = FFMatrixBlock>
inline pdb::Handle<Block> fetch_block(nb200_set *s, int br, int bc) {
    int tr, tc, rows, cols;
    nb200_set_info(s, &tr, &tc, &rows, &cols);
    pdb::Handle<Block> out = make_block<Block>(br, bc, rows, cols, tr, tc);
    if (nb200_download_block(s, br, bc, out->getRawDataHandle()->c_ptr(), rows, cols) != NB200_OK) {
        std::cerr << nb200_last_error() << std::endl;
        exit(1);
    }
    return out;
}

// one block of a row-sharded set, assembled from the devices that hold its rows
template <class Block = FFMatrixBlock>
inline pdb::Handle<Block> fetch_block_sharded(const std::string &db, const std::string &set, int br, int bc) {
    const int n = num_devices();
    std::vector<nb200_set *> sets(n, nullptr);
    std::vector<int> rows_of(n, 0);
    int tr = 0, tc = 0, R = 0, C = 0;
    for (int d = 0; d < n; d++) {
        die_on(nb200_use_device(d));
        die_on(nb200_set_lookup(db.c_str(), set.c_str(), &sets[d]));
        int r, c, rr, cc;
        die_on(nb200_set_info(sets[d], &r, &c, &rr, &cc));
        rows_of[d] = r; tr += r; tc = c; R = rr; C = cc;
    }
    nb200_use_device(0);
    pdb::Handle<Block> out = make_block<Block>(br, bc, R, C, tr, tc);   // allocated before any device work (zeroed)
    const int blo = br * R, bhi = std::min(blo + R, tr);
    int off = 0;
    for (int d = 0; d < n; d++) {
        const int lo = std::max(blo, off), hi = std::min(bhi, off + rows_of[d]);
        if (hi > lo) {
            die_on(nb200_use_device(d));
            die_on(nb200_download_rect(sets[d], lo - off, bc * C, out->getRawDataHandle()->c_ptr() + static_cast<size_t>(lo - blo) * C, hi - lo,
                                       C, C));
        }
        off += rows_of[d];
    }
    nb200_use_device(0);
    return out;
}

class FusedStage {
public:
    std::vector<int32_t> labels;
    nb200_ff_stats stats{};

    // several devices: the same lowering once per device, each on its own thread and its own (db, set) namespace
    bool run(const std::string &jobName, const std::vector<pdb::Handle<pdb::Computation>> &sinks, std::string &errMsg) {
        const int n = num_devices();
        if (n == 1) return run_one(jobName, sinks, errMsg);
        // outputs of a job that scans a sharded set are sharded themselves (every stage of the path is row-local)
        bool touches_sharded = false;
        std::vector<std::pair<std::string, std::string>> outs;
        for (auto &sink : sinks) {
            auto *w = dynamic_cast<pdb::WriteSetBase *>(sink.ptr().get());
            if (!w) { errMsg = "sink is not a WriteUserSet"; return false; }
            outs.push_back(std::make_pair(w->getDatabaseName(), w->getSetName()));
            touches_sharded |= scans_sharded(sink);
        }
        if (per_device.size() != static_cast<size_t>(n)) per_device.assign(n, nullptr);
        std::vector<std::string> errs(n);
        std::vector<char> oks(n, 1);
        std::vector<std::thread> th;
        for (int d = 0; d < n; d++)
            th.emplace_back([&, d] {
                if (nb200_use_device(d) != NB200_OK) { errs[d] = nb200_last_error(); oks[d] = 0; return; }
                if (!per_device[d]) per_device[d] = std::make_shared<FusedStage>();
                oks[d] = per_device[d]->run_one(jobName, sinks, errs[d]) ? 1 : 0;
            });
        for (auto &t : th) t.join();
        nb200_use_device(0);
        for (int d = 0; d < n; d++)
            if (!oks[d]) { errMsg = "device " + std::to_string(d) + ": " + errs[d]; return false; }
        if (touches_sharded) {
            for (auto &o : outs) sharded_sets().insert(o);
            labels.clear();
            for (int d = 0; d < n; d++) labels.insert(labels.end(), per_device[d]->labels.begin(), per_device[d]->labels.end());
        } else {
            labels = per_device[0]->labels;
        }
        stats = per_device[0]->stats;
        for (int d = 1; d < n; d++) if (per_device[d]->stats.ms_total > stats.ms_total) stats = per_device[d]->stats;
        return true;
    }

    bool run_one(const std::string &jobName, const std::vector<pdb::Handle<pdb::Computation>> &sinks, std::string &errMsg) {
        (void)jobName;
        try {
            for (auto &sink : sinks) {
                auto *w = dynamic_cast<pdb::WriteSetBase *>(sink.ptr().get());
                if (!w) throw std::runtime_error("sink is not a WriteUserSet");
                if (try_inference_unit(w)) continue;
                if (ConvStage::try_run(w)) continue;
                eval(w->getIthInput(0), w->getDatabaseName(), w->getSetName());
            }
        } catch (const std::exception &e) {
            errMsg = e.what();
            return false;
        }
        return true;
    }

private:
    int tmp_id = 0;
    std::map<std::string, std::string> fused_output_of;  // db:yo-set -> pending fused `output` set name
    std::vector<std::shared_ptr<FusedStage>> per_device;  // several devices: one lowering state per device

    static bool scans_sharded(const pdb::Handle<pdb::Computation> &c) {
        if (c == nullptr) return false;
        if (auto *scan = dynamic_cast<pdb::ScanSetBase *>(c.ptr().get()))
            return sharded_sets().count(std::make_pair(scan->getDatabaseName(), scan->getSetName())) != 0;
        for (int i = 0; i < c->getNumInputs(); i++)
            if (scans_sharded(c->getIthInput(i))) return true;
        return false;
    }

    static void check(int st) {
        if (st != NB200_OK) throw std::runtime_error(nb200_last_error());
    }
    static nb200_set *lookup(const std::string &db, const std::string &set) {
        nb200_set *s = nullptr;
        check(nb200_set_lookup(db.c_str(), set.c_str(), &s));
        return s;
    }
    template <class T> static T *as(const pdb::Handle<pdb::Computation> &c) { return dynamic_cast<T *>(c.ptr().get()); }
    std::string tmp() { return "__tmp" + std::to_string(tmp_id++); }
    static void dropout_must_be_zero(FFReluBiasSum *r) {
        if (r->dropout_rate == 0) return;
        // FFReluBiasSum.h:67,73-80 draws from srand(time(NULL)): irreproducible, so parity runs need 0
        if (getenv("NB200_DROPOUT_AS_ZERO")) {
            std::cerr << "warning: dropout_rate " << r->dropout_rate << " ignored (NB200_DROPOUT_AS_ZERO)" << std::endl;
            return;
        }
        throw std::runtime_error("FFReluBiasSum: dropout_rate must be 0 (the reference's dropout is time-seeded)");
    }

    // JoinComp(FFTransposeMult|FFInputLayerJoin) -> FFAggMatrix [-> bias join] as ONE device call
    nb200_set *lower_join(const pdb::Handle<pdb::Computation> &agg, int epilogue, nb200_set *bias, const std::string &db,
                          const std::string &out) {
        if (!as<FFAggMatrix>(agg)) throw std::runtime_error("expected FFAggMatrix below the bias join (no CPU fallback)");
        pdb::Handle<pdb::Computation> join = agg->getIthInput(0);
        int op;
        if (as<FFTransposeMult>(join)) op = NB200_OP_NT;
        else if (as<FFInputLayerJoin>(join) || as<LSTMHiddenState>(join)) op = NB200_OP_NN;
        else throw std::runtime_error("FFAggMatrix over an unsupported join (no CPU fallback)");
        nb200_set *a = eval(join->getIthInput(0), db, ""), *b = eval(join->getIthInput(1), db, "");
        if (as<LSTMHiddenState>(join)) {  // tanh of the second input (LSTMHiddenState.h:44-47), into a temporary: the
            nb200_set *t = nullptr;       // reference overwrites its input page, which is not reproduced
            check(nb200_sum_blocks(b, nullptr, nullptr, NB200_ACT_TANH, db.c_str(), tmp().c_str(), &t));
            b = t;
        }
        nb200_set *c = nullptr;
        check(nb200_gemm_blocks(op, a, b, epilogue, bias, precision_from_env(), db.c_str(), out.c_str(), &c));
        return c;
    }

    nb200_set *eval(const pdb::Handle<pdb::Computation> &c, const std::string &db, const std::string &out_name) {
        if (c == nullptr) throw std::runtime_error("computation has a missing input");
        const std::string out = out_name.empty() ? tmp() : out_name;
        if (auto *scan = as<pdb::ScanSetBase>(c)) return lookup(scan->getDatabaseName(), scan->getSetName());
        if (as<FFAggMatrix>(c)) return lower_join(c, NB200_EPI_NONE, nullptr, db, out);
        if (as<LASillyMultiply2Aggregate>(c)) {  // LA DSL: the same join + aggregation pair on MatrixBlock
            pdb::Handle<pdb::Computation> join = c->getIthInput(0);
            int op;
            if (as<LASillyMultiply1Join>(join)) op = NB200_OP_NN;
            else if (as<LASillyTransposeMultiply1Join>(join)) op = NB200_OP_TN;
            else throw std::runtime_error("LASillyMultiply2Aggregate over an unsupported join (no CPU fallback)");
            nb200_set *a = eval(join->getIthInput(0), db, ""), *b = eval(join->getIthInput(1), db, "");
            nb200_set *o = nullptr;
            check(nb200_gemm_blocks(op, a, b, NB200_EPI_NONE, nullptr, precision_from_env(), db.c_str(), out.c_str(), &o));
            return o;
        }
        if (auto *r = as<FFReluBiasSum>(c)) {
            dropout_must_be_zero(r);
            return lower_join(c->getIthInput(0), NB200_EPI_BIAS_RELU, eval(c->getIthInput(1), db, ""), db, out);
        }
        if (as<FFTransposeBiasSum>(c)) {
            nb200_set *z = lower_join(c->getIthInput(0), NB200_EPI_BIAS_TRANSPOSE, eval(c->getIthInput(1), db, ""), db,
                                      "__logits_of_" + out);
            nb200_set *view = nullptr;
            check(nb200_set_exp_view(z, db.c_str(), out.c_str(), &view));  // the reference's yo holds exp(.)
            return view;
        }
        if (auto *s3 = as<LSTMThreeWaySum>(c)) {
            nb200_set *x = eval(c->getIthInput(0), db, ""), *y = eval(c->getIthInput(1), db, ""), *z = eval(c->getIthInput(2), db, "");
            nb200_set *o = nullptr;
            const int act = s3->modifier == SumActivation::Sigmod ? NB200_ACT_SIGMOID
                            : s3->modifier == SumActivation::Tanh ? NB200_ACT_TANH : -1;  // "Invalid modifier!" (:92-94)
            check(nb200_sum_blocks(x, y, z, act, db.c_str(), out.c_str(), &o));
            return o;
        }
        if (as<LSTMTwoSum>(c)) {
            nb200_set *x = eval(c->getIthInput(0), db, ""), *y = eval(c->getIthInput(1), db, "");
            nb200_set *o = nullptr;
            check(nb200_sum_blocks(x, y, nullptr, NB200_ACT_NONE, db.c_str(), out.c_str(), &o));
            return o;
        }
        if ((as<FullyConnectedNetwork>(c) || as<SemanticClassifier>(c)) && num_devices() > 1 && scans_sharded(c->getIthInput(0)))
            // these UDFs reinterpret whole blocks; a set split over devices at row granularity has no whole blocks
            throw std::runtime_error(c->getComputationType() + ": a network UDF over a set sharded by NB200_DEVICES is not supported; "
                                     "take the set out of NB200_SHARD_SETS");
        if (auto *net = as<FullyConnectedNetwork>(c)) {
            nb200_set *x = eval(c->getIthInput(0), db, "");
            int br = 0;
            nb200_set_info(x, nullptr, nullptr, &br, nullptr);
            if (static_cast<uint32_t>(br) != net->sizeBatch)   // FullyConnectedNetwork.h:69-72 (the reference exits)
                throw std::runtime_error("batch size doesn't match: " + std::to_string(net->sizeBatch) + ", " + std::to_string(br));
            nb200_set *o = nullptr;
            check(nb200_udf_dense2(x, net->numFeatures, static_cast<int>(net->sizeBatch), net->numNeurons, net->numLabels, net->weights1.data(),
                                   net->bias1.data(), net->weights2.data(), net->bias2.data(), db.c_str(), out.c_str(), &o));
            return o;
        }
        if (auto *net = as<SemanticClassifier>(c)) {
            nb200_set *x = eval(c->getIthInput(0), db, "");
            int bc = 0;
            nb200_set_info(x, nullptr, nullptr, nullptr, &bc);   // sizeBatch = inNumCol (SemanticClassifier.h:57)
            const std::vector<double> w0(static_cast<size_t>(net->sizeDense0) * net->sizeEmbed, 1.0), b0(net->sizeDense0, 1.0),
                w1(static_cast<size_t>(net->sizeDense1) * net->sizeDense0, 1.0), b1(net->sizeDense1, 1.0);
            nb200_set *o = nullptr;
            check(nb200_udf_dense2(x, static_cast<int>(net->sizeEmbed), bc, static_cast<int>(net->sizeDense0), static_cast<int>(net->sizeDense1),
                                   w0.data(), b0.data(), w1.data(), b1.data(), db.c_str(), out.c_str(), &o));
            return o;
        }
        if (as<FFOutputLayer>(c)) {
            auto *ra = as<FFRowAggregate>(c->getIthInput(1));
            if (!ra) throw std::runtime_error("FFOutputLayer needs FFRowAggregate as its second input (no CPU fallback)");
            auto *s0 = as<pdb::ScanSetBase>(c->getIthInput(0));
            auto *s1 = as<pdb::ScanSetBase>(ra->getIthInput(0));
            if (s0 && s1 && s0->getSetName() == s1->getSetName()) {
                auto it = fused_output_of.find(s0->getDatabaseName() + ":" + s0->getSetName());
                if (it != fused_output_of.end()) {  // already produced by the fused stage: just publish it
                    nb200_set *o = lookup(s0->getDatabaseName(), it->second);
                    check(nb200_set_rename(o, db.c_str(), out.c_str()));
                    fused_output_of.erase(it);
                    return o;
                }
            }
            nb200_set *yo = eval(c->getIthInput(0), db, "");
            nb200_set *o = nullptr;
            check(nb200_softmax_rows(yo, db.c_str(), out.c_str(), &o, nullptr));
            return o;
        }
        throw std::runtime_error("computation graph not supported on the device (no CPU fallback): " + c->getComputationType());
    }

    // The graph of ff::inference_unit (SimpleFF.cc:340-389): Write(yo) <- FFTransposeBiasSum(FFAggMatrix(FFInputLayerJoin(
    // Scan(wo), FFReluBiasSum(FFAggMatrix(FFTransposeMult(Scan(w1), Scan(inputs))), Scan(b1)))), Scan(bo))
    bool try_inference_unit(pdb::WriteSetBase *w) {
        auto tbs = w->getIthInput(0);
        if (!as<FFTransposeBiasSum>(tbs)) return false;
        auto agg2 = tbs->getIthInput(0);
        auto *bo = as<pdb::ScanSetBase>(tbs->getIthInput(1));
        if (!as<FFAggMatrix>(agg2) || !bo) return false;
        auto j2 = agg2->getIthInput(0);
        if (!as<FFInputLayerJoin>(j2)) return false;
        auto *wo = as<pdb::ScanSetBase>(j2->getIthInput(0));
        auto relu = j2->getIthInput(1);
        auto *r = as<FFReluBiasSum>(relu);
        if (!wo || !r) return false;
        auto agg1 = relu->getIthInput(0);
        auto *b1 = as<pdb::ScanSetBase>(relu->getIthInput(1));
        if (!as<FFAggMatrix>(agg1) || !b1) return false;
        auto j1 = agg1->getIthInput(0);
        if (!as<FFTransposeMult>(j1)) return false;
        auto *w1 = as<pdb::ScanSetBase>(j1->getIthInput(0));
        auto *in = as<pdb::ScanSetBase>(j1->getIthInput(1));
        if (!w1 || !in) return false;
        dropout_must_be_zero(r);
        const std::string db = w->getDatabaseName();
        nb200_set *sin = lookup(in->getDatabaseName(), in->getSetName());
        int rows = 0;
        nb200_set_info(sin, &rows, nullptr, nullptr, nullptr);
        labels.assign(rows, -1);
        nb200_set *out = nullptr;
        check(nb200_ff_inference_unit(lookup(w1->getDatabaseName(), w1->getSetName()), lookup(wo->getDatabaseName(), wo->getSetName()),
                                      sin, lookup(b1->getDatabaseName(), b1->getSetName()),
                                      lookup(bo->getDatabaseName(), bo->getSetName()), precision_from_env(), db.c_str(),
                                      "__ff_output", &out, labels.data(), &stats));
        if (w->getSetName() != "yo") check(nb200_set_rename(lookup(db, "yo"), db.c_str(), w->getSetName().c_str()));
        fused_output_of[db + ":" + w->getSetName()] = "__ff_output";
        return true;
    }
};

}  // namespace nb200_host
