import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np
import test_closed_loop as T, _oracle as O
from gnss_sim_receiver_b200 import _hostapi as H
nsat=4; secs=float(sys.argv[1]) if len(sys.argv)>1 else 12.0
sc=T.e1_scenario(nsat)
bank=H.TrackingBank(T.CONF_E1,"Tracking_1B",nsat)
loops=[]
for k in range(nsat):
    dc = sc["pilot_code"][k] if len(sys.argv)>2 else sc["data_code"][k]; bank.set_codes(k, sc["pilot_code"][k], dc)
    a=(int(sc["prn"][k]),float(sc["acq_delay"][k]),float(sc["doppler"][k]+sc["acq_err"][k]),0,0)
    bank.start(k,*a)
    loops.append(O.RefLoop(T.e1_ref_params(), sc["pilot_code"][k], dc, *a))
got=[];ref=[[] for _ in range(nsat)]
for first,x in T.e1_chunks(sc,secs):
    bank.push(x); got.append(bank.step())
    for k in range(nsat): ref[k].append(loops[k].feed(x))
g=np.concatenate(got)
for k in range(nsat):
    gk=g[g["channel"]==k]; rk=np.concatenate(ref[k]); h=len(gk)//2
    for nm,a in (("gpu",gk),("ref",rk)):
        p=a["prompt_i"][h:]+1j*a["prompt_q"][h:]
        print(k,nm,"mean|I|=%.2f std(|I|)=%.2f meanQ=%.2f stdQ=%.2f cn0=%.3f dop=%.3f"%(np.mean(np.abs(p.real)),np.std(np.abs(p.real)),np.mean(p.imag*np.sign(p.real)),np.std(p.imag),a["cn0_db_hz"][-1],np.mean(a["carrier_doppler_hz"][h:])))
    d=gk["cn0_db_hz"]-rk["cn0_db_hz"]; print(k,"cn0 diff at 10%%,50%%,100%%: %.3f %.3f %.3f"%(d[len(d)//10],d[len(d)//2],d[-1]), "truth dop",sc["doppler"][k])
gk=g[g["channel"]==0]; rk=np.concatenate(ref[0])
print(gk.dtype.names)
for i in list(range(0,8))+[50,100,200,400,800,1200]:
    if i>=len(gk): break
    print(i,"gpu",[float("%.6g"%gk[nm][i]) for nm in ("prompt_i","prompt_q","carrier_doppler_hz","code_phase_samples","tracking_sample_counter","cn0_db_hz")])
    print(i,"ref",[float("%.6g"%rk[nm][i]) for nm in ("prompt_i","prompt_q","carrier_doppler_hz","code_phase_samples","tracking_sample_counter","cn0_db_hz")])
