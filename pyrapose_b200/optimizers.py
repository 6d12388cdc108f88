"""keras.optimizers.Adam as the reference configures it (PyraPose/bin/train.py:101: Adam(lr=lr, clipnorm=0.001)).

Keras 2.3.1 semantics: `clipnorm` clips by the GLOBAL norm over all gradients (optimizers.py get_gradients),
epsilon defaults to K.epsilon() = 1e-7, lr_t = lr * sqrt(1 - beta_2^t) / (1 - beta_1^t), no amsgrad / decay.
The update runs in csrc/train_kernels.cu (adam_update_kernel)."""


class Adam(object):
    def __init__(self, lr=0.001, beta_1=0.9, beta_2=0.999, epsilon=None, decay=0.0, amsgrad=False, clipnorm=0.0,
                 **kwargs):
        if 'learning_rate' in kwargs:
            lr = kwargs.pop('learning_rate')
        if kwargs:
            raise TypeError('Unexpected keyword argument passed to optimizer: ' + str(sorted(kwargs)))
        if amsgrad or decay:
            raise NotImplementedError('amsgrad / decay are not used by the reference training script')
        self.lr = float(lr)
        self.beta_1 = float(beta_1)
        self.beta_2 = float(beta_2)
        self.epsilon = 1e-7 if epsilon is None else float(epsilon)
        self.clipnorm = float(clipnorm)

    def get_config(self):
        return {'lr': self.lr, 'beta_1': self.beta_1, 'beta_2': self.beta_2, 'epsilon': self.epsilon,
                'clipnorm': self.clipnorm}
