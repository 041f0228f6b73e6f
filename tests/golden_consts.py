"""Constants shared by tests/golden/make_golden.py and the tests that replay its fixtures."""
LSTM_BEAM_END_BIAS = 0.35     # added to rnn.linear.bias[END] before the plain-LSTM beam fixtures were recorded
