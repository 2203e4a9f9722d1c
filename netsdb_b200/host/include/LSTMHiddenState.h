// LSTMHiddenState.h -- same header name as src/LSTM/headers/LSTMHiddenState.h; the class lives in lstm_udfs.h
#pragma once
#include "lstm_udfs.h"
