// LSTMTest -- one LSTM step over FFMatrixBlock sets, wired with the same Computation objects as the reference's driver
// (src/tests/source/LSTMTest.cc:204-543: FFInputLayerJoin + FFAggMatrix for every product, LSTMThreeWaySum for the
// gates, LSTMTwoSum for the cell state, LSTMHiddenState for the output), in database "lstm" with the reference's set
// names.  Unlike the reference driver (all-zero weights, no result check) the operands come from text files in the
// ff::load_matrix_data format and every result set is written back as text, so that the parity tests can compare
// with the oracle.
//
//   LSTMTest block dir      reads dir/<set>.txt for the 15 input sets, writes dir/<set>.out.txt for the 6 results
#include <cstdio>

#include "FFMatrixUtil.h"
#include "LSTMHiddenState.h"
#include "LSTMThreeWaySum.h"
#include "LSTMTwoSum.h"
#include "PDBClient.h"

using pdb::Computation;
using pdb::Handle;
using pdb::makeObject;

static pdb::PDBClient *client;
static std::string errMsg;

static Handle<Computation> scan(const char *set) { return makeObject<FFMatrixBlockScanner>("lstm", set); }
// FFAggMatrix(FFInputLayerJoin(a, b))
static Handle<Computation> product(Handle<Computation> a, Handle<Computation> b) {
    Handle<Computation> join = makeObject<FFInputLayerJoin>();
    join->setInput(0, a);
    join->setInput(1, b);
    Handle<Computation> agg = makeObject<FFAggMatrix>();
    agg->setInput(join);
    return agg;
}
static void run(Handle<Computation> top, const char *out_set) {
    Handle<Computation> writer = makeObject<FFMatrixWriter>("lstm", out_set);
    writer->setInput(top);
    if (!client->executeComputations(errMsg, writer)) {
        std::cout << "Computation failed. Message was: " << errMsg << "\n";
        exit(1);
    }
}
static void gate(const char *w, const char *u, const char *b, SumActivation act, const char *out_set) {
    Handle<Computation> sum = makeObject<LSTMThreeWaySum>(act);
    sum->setInput(0, product(scan(w), scan("x_t")));
    sum->setInput(1, product(scan(u), scan("h_t_1")));
    sum->setInput(2, scan(b));
    run(sum, out_set);
}
static void dump(const std::string &dir, const char *set) {
    int tr = 0, tc = 0;
    std::vector<double> dense;
    auto it = client->getSetIterator<FFMatrixBlock>("lstm", set);
    for (auto r : it) {
        if (dense.empty()) {
            tr = r->getTotalRowNums(); tc = r->getTotalColNums();
            dense.assign(static_cast<size_t>(tr) * tc, 0.0);
        }
        const double *d = r->getRawDataHandle()->c_ptr();
        for (int i = 0; i < r->getRowNums(); i++)
            for (int j = 0; j < r->getColNums(); j++) {
                const int row = r->getBlockRowIndex() * r->getRowNums() + i, col = r->getBlockColIndex() * r->getColNums() + j;
                if (row < tr && col < tc) dense[static_cast<size_t>(row) * tc + col] = d[static_cast<size_t>(i) * r->getColNums() + j];
            }
    }
    FILE *f = fopen((dir + "/" + set + ".out.txt").c_str(), "w");
    fprintf(f, "%d %d\n", tr, tc);
    for (int i = 0; i < tr; i++)
        for (int j = 0; j < tc; j++) fprintf(f, "%.17g%c", dense[static_cast<size_t>(i) * tc + j], j + 1 == tc ? '\n' : ' ');
    fclose(f);
}

int main(int argc, char *argv[]) {
    if (argc < 3) {
        std::cout << "Usage: block dir" << std::endl;
        return -1;
    }
    const int block = atoi(argv[1]);
    const std::string dir = argv[2];
    pdb::PDBLoggerPtr clientLogger = std::make_shared<pdb::PDBLogger>("LSTMclientLog");
    pdb::PDBClient pdbClient(8108, "localhost", clientLogger, false, true);
    client = &pdbClient;
    pdbClient.createDatabase("lstm", errMsg);
    for (const char *s : {"w_f", "u_f", "b_f", "w_i", "u_i", "b_i", "w_o", "u_o", "b_o", "w_c", "u_c", "b_c", "x_t", "h_t_1", "c_t_1"}) {
        pdbClient.createSet<FFMatrixBlock>("lstm", s, errMsg, static_cast<size_t>(64) << 20, s);
        ff::load_matrix_data(pdbClient, dir + "/" + s + ".txt", "lstm", s, block, block, false, false, errMsg);
    }
    gate("w_f", "u_f", "b_f", SumActivation::Sigmod, "f_t");
    gate("w_i", "u_i", "b_i", SumActivation::Sigmod, "i_t");
    gate("w_o", "u_o", "b_o", SumActivation::Sigmod, "o_t");
    gate("w_c", "u_c", "b_c", SumActivation::Tanh, "c_t_temp");
    {   // c_t = f_t . c_t_1 + i_t . c_t_temp  (matrix products, as the reference driver wires them)
        Handle<Computation> sum = makeObject<LSTMTwoSum>();
        sum->setInput(0, product(scan("f_t"), scan("c_t_1")));
        sum->setInput(1, product(scan("i_t"), scan("c_t_temp")));
        run(sum, "c_t");
    }
    {   // h_t = o_t . tanh(c_t)
        Handle<Computation> join = makeObject<LSTMHiddenState>();
        join->setInput(0, scan("o_t"));
        join->setInput(1, scan("c_t"));
        Handle<Computation> agg = makeObject<FFAggMatrix>();
        agg->setInput(join);
        run(agg, "h_t");
    }
    for (const char *s : {"f_t", "i_t", "o_t", "c_t_temp", "c_t", "h_t"}) dump(dir, s);
    return 0;
}
