import sys; sys.path.insert(0,'/root/repo')
from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_weights
for v in "123":
    h=HParams.builtin(v); g=Generator(h).load_weights(synthetic_weights(h,"A"))
    f=g.layer_folds(); e=g.layer_engines()
    print(v, {k:(f[k],e[k]) for k in f if f[k][0]!=1 or f[k][1]!=1 or e[k]!='tc'})
