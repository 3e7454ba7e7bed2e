"""dlas -- B200-native drop-in for the training hot path of OmarZOS/deep-learning-at-scale.

Public surface mirrors the Keras names the reference scripts use (cnn.py:10-15, :110-132)."""
from ._lib import (DlasError, GILBERT_MUFU, GILBERT_POLY, LOSS_MAE_CLIP, LOSS_MSE, MODE_FP32, MODE_TC,
                   OPT_ADAM, OPT_SGD, load)
from .models import SGD, Adam, Dense, EarlyStopping, ModelCheckpoint, Sequential, mae_clip
from .ops import LagWindows, gilbert
from .recurrent import LSTM, LSTMRegressor
from .convolutional import CnnRegressor, Convolution1D, Dropout, Flatten

__all__ = ["Sequential", "Dense", "SGD", "Adam", "EarlyStopping", "ModelCheckpoint", "mae_clip", "gilbert", "LagWindows", "LSTM", "LSTMRegressor",
           "Convolution1D", "Dropout", "Flatten", "CnnRegressor",
           "load", "DlasError"]
