import sys, os, logging
sys.path.insert(0, os.getcwd())
import numpy as np
from ocrd_keraslm_b200 import synth
from ocrd_keraslm_b200.lib import Rater
log = logging.getLogger('p'); log.setLevel(logging.CRITICAL)
wd = synth.weights(7, width=128, depth=2, voc_size=256, style='trained')
r = Rater.from_weights(wd, synth.vocabulary(), 256, logger=log)
text = synth.text(5, 20000)
import time
r.rate(text, [3]); 
t0 = time.time(); p = r.rate(text, [3]); dt = time.time() - t0
i = r._native.info()
print('cl_cluster', i['cl_cluster'], 'cl_launches', i['cl_launches'], 'launches', i['kernel_launches'], 'chars/s %.0f' % (len(text) / dt), 'kernel ms', i['last_kernel_ms'])
