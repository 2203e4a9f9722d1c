// LSTMTwoSum.h -- same header name as src/LSTM/headers/LSTMTwoSum.h; the class lives in lstm_udfs.h
#pragma once
#include "lstm_udfs.h"
