// net_udfs.h -- the two SelectionComps of the reference that hold a whole two-layer network, by their netsDB names,
// base class and constructor signatures:
//   FullyConnectedNetwork   src/FF_proj/headers/FullyConnectedNetwork.h:17-124   (driver: src/tests/source/FCProjTest.cc)
//   SemanticClassifier      src/word2vec/headers/SemanticClassifier.h:16-143     (driver: src/word2vec/source/TestSemanticClassifier.cc)
// The projection is not run per block on the host: ff_stage.h lowers the computation to nb200_udf_dense2, which pushes
// every block of the input set through two fp64 tensor-core GEMMs at once.
#pragma once
#include <random>

#include "FFMatrixBlock.h"
#include "ff_udfs.h"

using namespace pdb;

// weights1: numNeurons x numFeatures, weights2: numLabels x numNeurons, row-major (FullyConnectedNetwork.h:79-81,96-97).
// The reference fills them in the constructor from std::random_device with magnitudes in [1e-4, 0.5] and a random sign
// (FullyConnectedNetwork.cc:26-41), so two runs of its driver never agree; here NB200_FC_SEED makes a run repeatable.
class FullyConnectedNetwork : public SelectionComp<FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FullyConnectedNetwork() {}
    FullyConnectedNetwork(int numFeatures, int numLabels, int numNeurons, uint32_t sizeBatch)
        : numFeatures(numFeatures), numLabels(numLabels), numNeurons(numNeurons), sizeBatch(sizeBatch) {
        const char *seed = getenv("NB200_FC_SEED");
        std::mt19937_64 gen(seed ? strtoull(seed, nullptr, 10) : std::random_device{}());
        auto draw = [&gen](std::vector<double> &v, size_t n) {
            std::uniform_real_distribution<double> mag(0.0001, 0.5);
            v.resize(n);
            for (double &x : v) x = (gen() & 1) ? -mag(gen) : mag(gen);
        };
        draw(weights1, static_cast<size_t>(numFeatures) * numNeurons);
        draw(bias1, numNeurons);
        draw(weights2, static_cast<size_t>(numNeurons) * numLabels);
        draw(bias2, numLabels);
    }
    Lambda<bool> getSelection(Handle<FFMatrixBlock> checkMe) override {
        return makeLambda(checkMe, [](Handle<FFMatrixBlock> &) { return true; });
    }
    Lambda<Handle<FFMatrixBlock>> getProjection(Handle<FFMatrixBlock> in) override {
        return makeLambda(in, [](Handle<FFMatrixBlock> &) {
            nb200_host::no_host_projection("FullyConnectedNetwork");
            return Handle<FFMatrixBlock>();
        });
    }
    int numFeatures = 0, numLabels = 0, numNeurons = 0;
    uint32_t sizeBatch = 0;
    std::vector<double> weights1, bias1, weights2, bias2;
};

// dense 0: sizeDense0 x sizeEmbed, dense 1: sizeDense1 x sizeDense0; the projection builds every weight and bias as 1
// (SemanticClassifier.h:66-72) and takes the batch from the block's column count (:57).
class SemanticClassifier : public SelectionComp<FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    SemanticClassifier() : sizeEmbed(500), sizeDense0(16), sizeDense1(1) {}
    SemanticClassifier(uint32_t sizeEmbed, uint32_t sizeDense0, uint32_t sizeDense1)
        : sizeEmbed(sizeEmbed), sizeDense0(sizeDense0), sizeDense1(sizeDense1) {}
    Lambda<bool> getSelection(Handle<FFMatrixBlock> checkMe) override {
        return makeLambda(checkMe, [](Handle<FFMatrixBlock> &) { return true; });
    }
    Lambda<Handle<FFMatrixBlock>> getProjection(Handle<FFMatrixBlock> in) override {
        return makeLambda(in, [](Handle<FFMatrixBlock> &) {
            nb200_host::no_host_projection("SemanticClassifier");
            return Handle<FFMatrixBlock>();
        });
    }
    uint32_t sizeEmbed, sizeDense0, sizeDense1;
};
