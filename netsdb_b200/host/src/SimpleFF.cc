// ff:: driver functions over the shim (counterpart of src/FF/source/SimpleFF.cc).  inference_unit builds
// the same two Computation graphs as the reference (:340-389 and :398-415) and submits them with
// executeComputations, which lowers them to the device.
#include "SimpleFF.h"

#include <chrono>

#include "PDBClient.h"

namespace ff {

void loadLibrary(pdb::PDBClient &pdbClient, std::string path) {
    std::string errMsg;
    if (!pdbClient.registerType(path, errMsg)) {
        std::cout << "Couldnt include " << path << ": " << errMsg << std::endl;
        exit(-1);
    }
}

void createSet(pdb::PDBClient &pdbClient, std::string dbName, std::string setName, std::string setName1, int size) {
    std::string errMsg;
    pdbClient.removeSet(dbName, setName, errMsg);
    if (!pdbClient.createSet<FFMatrixBlock>(dbName, setName, errMsg, (size_t)size * 1024 * 1024, setName1)) {
        std::cout << "Not able to create set " + setName + ": " + errMsg;
        exit(-1);
    }
}

void createDatabase(pdb::PDBClient &pdbClient, std::string dbName) {
    std::string errMsg;
    if (!pdbClient.createDatabase(dbName, errMsg)) {
        std::cout << "Not able to create database: " << errMsg << std::endl;
        exit(-1);
    }
}

void setup(pdb::PDBClient &pdbClient, std::string) {
    // the reference registers one shared library per UDF here (SimpleFF.cc:69-82); the UDFs are compiled in
    for (const char *lib : {"libFFMatrixMeta.so", "libFFMatrixData.so", "libFFMatrixBlock.so", "libFFMatrixBlockScanner.so",
                            "libFFMatrixWriter.so", "libFFInputLayerJoin.so", "libFFReluBiasSum.so", "libFFTransposeMult.so",
                            "libFFTransposeBiasSum.so", "libFFRowAggregate.so", "libFFOutputLayer.so", "libFFAggMatrix.so"})
        loadLibrary(pdbClient, std::string("libraries/") + lib);
}

void cleanup(pdb::PDBClient &pdbClient, std::string database) {
    std::string errMsg;
    for (const char *s : {"inputs", "label", "w1", "b1", "wo", "bo", "y1", "yo", "output"}) pdbClient.removeSet(database, s, errMsg);
}

// One layer of the 3-layer pipeline as its own graph + executeComputations, as in the reference
// (SimpleFF.cc:99-147 layer 1 with FFTransposeMult; :166-214, :231-275 the FFInputLayerJoin layers).
static void run_layer(pdb::PDBClient &pdbClient, const std::string &database, bool first, const std::string &w,
                      const std::string &in, const std::string &bias, bool last, double dropout_rate, const std::string &out,
                      const char *jobName) {
    using pdb::Computation;
    using pdb::Handle;
    using pdb::makeObject;
    std::string errMsg;
    const pdb::UseTemporaryAllocationBlock tempBlock{1024 * 1024 * 128};
    Handle<Computation> readW = makeObject<FFMatrixBlockScanner>(database, w);
    Handle<Computation> readIn = makeObject<FFMatrixBlockScanner>(database, in);
    Handle<Computation> join;
    if (first) join = makeObject<FFTransposeMult>();
    else join = makeObject<FFInputLayerJoin>();
    join->setInput(0, readW);
    join->setInput(1, readIn);
    Handle<Computation> agg = makeObject<FFAggMatrix>();
    agg->setInput(join);
    Handle<Computation> readB = makeObject<FFMatrixBlockScanner>(database, bias);
    Handle<Computation> act;
    if (last) act = makeObject<FFTransposeBiasSum>();
    else act = makeObject<FFReluBiasSum>(dropout_rate);
    act->setInput(0, agg);
    act->setInput(1, readB);
    Handle<Computation> writer = makeObject<FFMatrixWriter>(database, out);
    writer->setInput(act);
    auto t0 = std::chrono::high_resolution_clock::now();
    if (!pdbClient.executeComputations(errMsg, jobName, false, writer)) {
        std::cout << "Computation failed. Message was: " << errMsg << "\n";
        exit(1);
    }
    std::cout << jobName << " Stage Time Duration: "
              << std::chrono::duration<float>(std::chrono::high_resolution_clock::now() - t0).count() << " secs." << std::endl;
}

void inference_compute(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string w2, std::string wo,
                       std::string inputs, std::string b1, std::string b2, std::string bo, double dropout_rate, bool) {
    run_layer(pdbClient, database, true, w1, inputs, b1, false, dropout_rate, "y1", "inference-1");
    run_layer(pdbClient, database, false, w2, "y1", b2, false, dropout_rate, "y2", "inference-2");
    run_layer(pdbClient, database, false, wo, "y2", bo, true, dropout_rate, "yo", "inference-3");
}

void inference(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string w2, std::string wo, std::string inputs,
               std::string b1, std::string b2, std::string bo, std::string output, double dropout_rate, bool enablePartition) {
    using pdb::Computation;
    using pdb::Handle;
    using pdb::makeObject;
    inference_compute(pdbClient, database, w1, w2, wo, inputs, b1, b2, bo, dropout_rate, enablePartition);
    std::string errMsg;
    const pdb::UseTemporaryAllocationBlock tempBlock{1024 * 1024 * 128};
    Handle<Computation> readYo = makeObject<FFMatrixBlockScanner>(database, "yo");
    Handle<Computation> expSum = makeObject<FFRowAggregate>();
    expSum->setInput(readYo);
    Handle<Computation> softmax = makeObject<FFOutputLayer>();
    softmax->setInput(0, readYo);
    softmax->setInput(1, expSum);
    Handle<Computation> writeOut = makeObject<FFMatrixWriter>(database, output);
    writeOut->setInput(softmax);
    if (!pdbClient.executeComputations(errMsg, "inference-3", false, writeOut)) {
        std::cout << "Computation failed. Message was: " << errMsg << "\n";
        exit(1);
    }
}

void inference_unit(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string wo, std::string inputs,
                    std::string b1, std::string bo, std::string output, double dropout_rate, bool) {
    using pdb::Computation;
    using pdb::Handle;
    using pdb::makeObject;
    std::string errMsg;
    const pdb::UseTemporaryAllocationBlock tempBlock{1024 * 1024 * 128};

    // y1 = relu(w1 . inputs^T + b1)
    Handle<Computation> readW1 = makeObject<FFMatrixBlockScanner>(database, w1);
    Handle<Computation> readX = makeObject<FFMatrixBlockScanner>(database, inputs);
    Handle<Computation> mult1 = makeObject<FFTransposeMult>();
    mult1->setInput(0, readW1);
    mult1->setInput(1, readX);
    Handle<Computation> agg1 = makeObject<FFAggMatrix>();
    agg1->setInput(mult1);
    Handle<Computation> readB1 = makeObject<FFMatrixBlockScanner>(database, b1);
    Handle<Computation> relu = makeObject<FFReluBiasSum>(dropout_rate);
    relu->setInput(0, agg1);
    relu->setInput(1, readB1);
    // yo = exp((wo . y1 + bo)^T)
    Handle<Computation> readWo = makeObject<FFMatrixBlockScanner>(database, wo);
    Handle<Computation> mult2 = makeObject<FFInputLayerJoin>();
    mult2->setInput(0, readWo);
    mult2->setInput(1, relu);
    Handle<Computation> agg2 = makeObject<FFAggMatrix>();
    agg2->setInput(mult2);
    Handle<Computation> readBo = makeObject<FFMatrixBlockScanner>(database, bo);
    Handle<Computation> biasT = makeObject<FFTransposeBiasSum>();
    biasT->setInput(0, agg2);
    biasT->setInput(1, readBo);
    Handle<Computation> writeYo = makeObject<FFMatrixWriter>(database, "yo");
    writeYo->setInput(biasT);

    auto t0 = std::chrono::high_resolution_clock::now();
    if (!pdbClient.executeComputations(errMsg, "inference-unit-intermediate", false, writeYo)) {
        std::cout << "Computation failed. Message was: " << errMsg << "\n";
        exit(1);
    }
    auto t1 = std::chrono::high_resolution_clock::now();
    std::cout << "Inference-unit Intermediate Stage Time Duration: " << std::chrono::duration<float>(t1 - t0).count() << " secs."
              << std::endl;

    // output = yo / rowsum(yo)
    Handle<Computation> readYo = makeObject<FFMatrixBlockScanner>(database, "yo");
    Handle<Computation> expSum = makeObject<FFRowAggregate>();
    expSum->setInput(readYo);
    Handle<Computation> softmax = makeObject<FFOutputLayer>();
    softmax->setInput(0, readYo);
    softmax->setInput(1, expSum);
    Handle<Computation> writeOut = makeObject<FFMatrixWriter>(database, output);
    writeOut->setInput(softmax);
    t0 = std::chrono::high_resolution_clock::now();
    if (!pdbClient.executeComputations(errMsg, "inference-unit", false, writeOut)) {
        std::cout << "Computation failed. Message was: " << errMsg << "\n";
        exit(1);
    }
    t1 = std::chrono::high_resolution_clock::now();
    std::cout << "Inference-unit Stage Time Duration: " << std::chrono::duration<float>(t1 - t0).count() << " secs." << std::endl;
}

}  // namespace ff
