for r in 1 2; do
  echo "NT default run $r"; python scripts/dsweep_profile.py c1,c3 device 5 | grep -v '"rep": [01],' | cut -c1-90
  echo "NT 512 run $r"; TCI_DSWEEP_NT=512 python scripts/dsweep_profile.py c1,c3 device 5 | grep -v '"rep": [01],' | cut -c1-90
done
